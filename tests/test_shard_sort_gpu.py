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


@pytest.mark.parametrize("kind", ["uniform", "few", "one_bucket", "clustered"])
@pytest.mark.parametrize("n,world", [(3, 2), (1000, 3), (8192 * 3 + 5, 2), (100003, 8), ((1 << 20) + 77, 4)])
@pytest.mark.parametrize("f64", [False, True])
def test_exchange_on_one_device(cuda_env, kind, n, world, f64):
    """col_shard_count / col_shard_push / col_shard_sort_received with every "rank" played on one device
    (the receive buffers of all ranks are local allocations): each rank ends up with its positions of the
    full stable sort -- codes, ids -- and can reach the coordinates and radius of every leaf through
    place_sorted."""
    import ctypes as C
    import torch
    from collision_b200 import _lib
    dev, cq = cuda_env
    L, P, S = _lib.lib(), _lib.ptr, cq.cuda_stream
    fdt, isz, code = (torch.float64, 8, 1) if f64 else (torch.float32, 4, 0)
    if f64 and n > 200000:
        pytest.skip("covered in f32")
    codes = _codes(kind, n, dev, n % 1000 + len(kind) + world)
    ids = torch.arange(n, dtype=torch.int32, device=dev)
    g = torch.Generator(device=dev).manual_seed(3)
    coords = torch.rand(n, 4, dtype=fdt, device=dev, generator=g)
    radii = torch.rand(n, dtype=fdt, device=dev, generator=g)
    want_c, want_i = torch.empty_like(codes), torch.empty_like(ids)
    cq.wait_stream(torch.cuda.current_stream(dev))
    assert L.col_sort_pairs(P(codes), P(want_c), P(ids), P(want_i), n, 4, 4, 30,
                            P(_lib.alloc_bytes(dev, L.col_sort_scratch_bytes(n, 4, 4))), S) == 0
    cuts = [n * r // world for r in range(world + 1)]          # slices of the input AND ranges of positions
    pos = []
    for r in range(world):
        pos += [max(cuts[r], 1) - 1, min(cuts[r + 1], n - 1)] if cuts[r + 1] > cuts[r] else [1, 0]
    pos_t = torch.tensor(pos, dtype=torch.int32, device=dev)
    hist_all = torch.zeros(world * 256, dtype=torch.int32, device=dev)
    # every slice's codes in a 16-byte aligned buffer of its own, like a rank's
    own = [codes[cuts[r]:cuts[r + 1]].clone() for r in range(world)]
    for r in range(world):
        assert L.col_shard_count(P(own[r]) if own[r].numel() else None, own[r].numel(), 30,
                                 P(hist_all) + 1024 * r, S) == 0
    recv = [[torch.full((n * w,), -1, dtype=torch.int32, device=dev) for w in (1, 1, isz, isz // 4)] for _ in range(world)]
    tables = [(C.c_uint64 * world)(*[recv[r][j].data_ptr() for r in range(world)]) for j in range(4)]
    plans = [torch.zeros(8 * world, dtype=torch.int32, device=dev) for _ in range(world)]
    cq.wait_stream(torch.cuda.current_stream(dev))
    for r in range(world):
        b, e = cuts[r], cuts[r + 1]
        scratch = _lib.alloc_bytes(dev, L.col_shard_push_scratch_bytes(e - b, world))
        oc, orad = (coords[b:e].contiguous(), radii[b:e].contiguous())
        assert L.col_shard_push(P(hist_all), world, r, P(pos_t), n, P(own[r]) if e > b else None, e - b, b,
                                P(oc) if e > b else None, P(orad) if e > b else None, code, 30, *tables, P(plans[r]),
                                P(scratch), S) == 0
        cq.synchronize()
    plan0 = plans[0].view(world, 8)
    for r in range(1, world):      # every rank derives the same plan (but for its own base column)
        assert torch.equal(plans[r].view(world, 8)[:, :4], plan0[:, :4])
    for r in range(world):
        if cuts[r + 1] == cuts[r]:
            assert plan0[r].tolist()[:4] == [0, 0, 256, 256]
            continue
        off, cnt, d_lo, d_hi = plan0[r].tolist()[:4]
        assert off <= pos[2 * r] and pos[2 * r + 1] < off + cnt <= n
        rc, ri, rco, rra = recv[r]
        assert bool((rc[cnt:] == -1).all()) and bool((ri[cnt:] == -1).all())       # nothing beyond the count
        got_ids = ri[:cnt].to(torch.int64)
        assert bool((got_ids[1:] > got_ids[:-1]).all())                              # input order
        assert torch.equal(rc[:cnt], codes[got_ids])
        assert torch.equal(rco.view(fdt)[:cnt * 4].view(cnt, 4), coords[got_ids])
        assert torch.equal(rra.view(fdt)[:cnt], radii[got_ids])
        out_c = torch.full((n,), SENTINEL, dtype=torch.int32, device=dev)
        out_p = torch.full((n,), SENTINEL, dtype=torch.int32, device=dev)
        out_i = torch.full((n,), SENTINEL, dtype=torch.int32, device=dev)
        scratch = _lib.alloc_bytes(dev, L.col_shard_sort_received_scratch_bytes(n))
        cq.wait_stream(torch.cuda.current_stream(dev))
        assert L.col_shard_sort_received(P(rc), P(ri), n, P(plans[r]) + 32 * r, 30, P(out_c), P(out_p), P(out_i),
                                         P(scratch), S) == 0
        cq.synchronize()
        assert torch.equal(out_c[off:off + cnt], want_c[off:off + cnt])
        assert torch.equal(out_i[off:off + cnt], want_i[off:off + cnt])
        assert torch.equal(ri[:cnt][out_p[off:off + cnt].to(torch.int64)], want_i[off:off + cnt])   # place -> id
        for t in (out_c, out_p, out_i):
            assert bool((t[:off] == SENTINEL).all()) and bool((t[off + cnt:] == SENTINEL).all())
