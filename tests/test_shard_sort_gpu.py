"""col_shard_sort (the sort dealt out between the ranks of mode "sharded_build"): the selected part of
the sorted order must be bit-identical to the same positions of the full stable sort
(col_sort_pairs with ids = 0..n-1), and nothing outside it may be written."""
import pytest

pytestmark = pytest.mark.gpu

SENTINEL = -559038737   # 0xDEADBEEF as int32


def _codes(kind, n, dev, seed):
    import torch
    g = torch.Generator(device=dev).manual_seed(seed)
    if kind == "uniform":
        return torch.randint(0, 1 << 30, (n,), dtype=torch.int32, device=dev, generator=g)
    if kind == "few":          # many duplicates: stability decides the ids
        return torch.randint(0, 7, (n,), dtype=torch.int32, device=dev, generator=g) * 0x01234567 % (1 << 30)
    if kind == "one_bucket":   # every code in the same top digit: degrades to the full sort
        return (5 << 22) | torch.randint(0, 1 << 22, (n,), dtype=torch.int32, device=dev, generator=g)
    if kind == "sorted":
        return (torch.arange(n, dtype=torch.int64, device=dev) * ((1 << 30) // max(n, 1))).to(torch.int32)
    if kind == "clustered":    # three narrow clusters of codes
        c = torch.randint(0, 3, (n,), dtype=torch.int32, device=dev, generator=g)
        return c * (300 << 20) + torch.randint(0, 1 << 21, (n,), dtype=torch.int32, device=dev, generator=g)
    raise ValueError(kind)


@pytest.mark.parametrize("kind", ["uniform", "few", "one_bucket", "sorted", "clustered"])
@pytest.mark.parametrize("n", [1, 2, 255, 8192, 8193, 100003, (1 << 21) + 77])
def test_shard_sort_matches_full_sort(cuda_env, kind, n):
    import torch
    from collision_b200 import _lib
    dev, cq = cuda_env
    L, P, S = _lib.lib(), _lib.ptr, cq.cuda_stream
    codes = _codes(kind, n, dev, n % 1000 + len(kind))
    ids = torch.arange(n, dtype=torch.int32, device=dev)
    want_c, want_i = torch.empty_like(codes), torch.empty_like(ids)
    full_scratch = _lib.alloc_bytes(dev, L.col_sort_scratch_bytes(n, 4, 4))
    scratch = _lib.alloc_bytes(dev, L.col_shard_sort_scratch_bytes(n))
    rng = torch.zeros(4, dtype=torch.int32, device=dev)
    cq.wait_stream(torch.cuda.current_stream(dev))
    assert L.col_sort_pairs(P(codes), P(want_c), P(ids), P(want_i), n, 4, 4, 30, P(full_scratch), S) == 0
    cq.synchronize()
    # the same thing from torch, once (stable sort on the codes)
    order = torch.sort(codes.to(torch.int64), stable=True).indices
    assert torch.equal(want_i.to(torch.int64), order)
    world = 4
    cuts = [n * r // world for r in range(world + 1)]
    for r in range(world):
        lo, hi = max(cuts[r], 1) - 1, min(cuts[r + 1], n - 1)
        if cuts[r + 1] == cuts[r]:
            continue
        got_c = torch.full((n,), SENTINEL, dtype=torch.int32, device=dev)
        got_i = torch.full((n,), SENTINEL, dtype=torch.int32, device=dev)
        cq.wait_stream(torch.cuda.current_stream(dev))
        for _ in range(2):      # the scratch is reusable as it is left
            assert L.col_shard_sort(P(codes), n, lo, hi, 30, P(got_c), P(got_i), P(rng), P(scratch), S) == 0
        cq.synchronize()
        off, cnt, d_lo, d_hi = rng.tolist()
        assert off <= lo and hi < off + cnt <= n and 0 <= d_lo <= d_hi < 256
        assert torch.equal(got_c[off:off + cnt], want_c[off:off + cnt])
        assert torch.equal(got_i[off:off + cnt], want_i[off:off + cnt])
        # nothing outside the selected positions is touched
        assert bool((got_c[:off] == SENTINEL).all()) and bool((got_c[off + cnt:] == SENTINEL).all())
        assert bool((got_i[:off] == SENTINEL).all()) and bool((got_i[off + cnt:] == SENTINEL).all())
        if kind == "uniform" and n >= 100003:
            assert cnt < n * 0.3        # a quarter of the positions + at most a bucket either side


def test_shard_sort_arguments(cuda_env):
    import torch
    from collision_b200 import _lib
    dev, cq = cuda_env
    L, P, S = _lib.lib(), _lib.ptr, cq.cuda_stream
    n = 1000
    codes = torch.zeros(n + 1, dtype=torch.int32, device=dev)
    out = torch.zeros(n, dtype=torch.int32, device=dev)
    rng = torch.zeros(4, dtype=torch.int32, device=dev)
    scratch = _lib.alloc_bytes(dev, L.col_shard_sort_scratch_bytes(n))
    assert L.col_shard_sort(P(codes), n, 5, 4, 30, P(out), P(out), P(rng), P(scratch), S) == -1      # lo > hi
    assert L.col_shard_sort(P(codes), n, n, n, 30, P(out), P(out), P(rng), P(scratch), S) == -1      # beyond the end
    assert L.col_shard_sort(P(codes) + 4, n, 0, 4, 30, P(out), P(out), P(rng), P(scratch), S) == -1  # misaligned codes
    assert L.col_shard_sort(P(codes), 0, 0, 0, 30, P(out), P(out), P(rng), P(scratch), S) == -1
    cq.synchronize()
