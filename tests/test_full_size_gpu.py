"""The configurations of BASELINE.json at their FULL sizes.

Where the CPU oracle finishes in under a minute the comparison is bit-exact against it
(configs[1]: 2^20 uniform spheres; configs[3]: 2^24 spheres in Gaussian clusters -- north_star's
"full-pipeline 16M-sphere collision detection bit-exact").  Beyond that the checks are the
size-independent properties of the domain, computed with torch on the device as the checker:
sortedness / permutation / stability of the 2^26-key sort, the running sum of the scan, and for
2^27 spheres: the list is grouped by query in sorted order, every row pairs a query with a LATER
leaf whose box really overlaps, and its length equals two independent counts (the count-only
walk on the traversal records and the reference-layout kernel on the Node/Bound arrays).
All calls go through the C ABI."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def uniform_scene(n, seed=4):
    rs = np.random.RandomState(seed)
    c = np.zeros((n, 4), np.float32)
    c[:, :3] = rs.random_sample((n, 3))
    return c, np.full(n, 0.5 * n ** (-1.0 / 3.0), np.float32)


def clustered_scene(n, seed=4, k=1024, sigma=0.02):
    rs = np.random.RandomState(seed)
    centres = rs.random_sample((k, 3)).astype(np.float32)
    which = rs.randint(0, k, n)
    c = np.zeros((n, 4), np.float32)
    c[:, :3] = centres[which] + rs.standard_normal((n, 3)).astype(np.float32) * np.float32(sigma)
    return c, np.full(n, 0.5 * n ** (-1.0 / 3.0), np.float32)


def collide_on_device(cuda_env, c, r):
    """count-only call, then the list with exactly that capacity (the reference's usage)"""
    import torch
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    n = len(r)
    col = Collider(dev, n, 8, 128)
    coords, radii = torch.from_numpy(c).to(dev), torch.from_numpy(r).to(dev)
    n_buf = torch.zeros(1, dtype=torch.int32, device=dev)
    col.get_collisions(cq, coords, radii, n_buf, None, 0).synchronize()
    total = int(n_buf.item()) & 0xFFFFFFFF
    pairs = torch.empty(max(total, 1) * 2, dtype=torch.int32, device=dev)
    col.get_collisions(cq, coords, radii, n_buf, pairs, total).synchronize()
    assert int(n_buf.item()) & 0xFFFFFFFF == total
    return col, coords, radii, total, pairs.view(-1, 2)[:total]


@pytest.mark.parametrize("n,scene", [(1 << 20, uniform_scene), (1 << 24, clustered_scene)],
                         ids=["configs1-1M-uniform", "configs3-16M-clustered"])
def test_full_pipeline_bit_exact(cuda_env, oracle, n, scene):
    c, r = scene(n)
    want = oracle.collide(c, r)
    col, _, _, total, pairs = collide_on_device(cuda_env, c, r)
    assert total == want["total"]
    got = pairs.cpu().numpy().view(np.uint32)
    assert np.array_equal(got, want["pairs"])
    # the stages behind it: sorted codes + ids, Node array
    assert np.array_equal(col._codes_bufs[1][:4 * n].cpu().numpy().view(np.uint32), want["sorted_codes"])
    assert np.array_equal(col._ids_bufs[1][:4 * n].cpu().numpy().view(np.uint32), want["sorted_ids"])
    nodes = col._nodes_buf[:16 * (2 * n - 1)].cpu().numpy().view(np.uint32).reshape(-1, 4)
    ref = want["nodes"].view(np.uint32).reshape(-1, 4)
    assert np.array_equal(nodes[:, 1:3], ref[:, 1:3])                 # right_edge, data[0]
    assert np.array_equal(nodes[1:, 0], ref[1:, 0])                   # parent (the root's is never written upstream)
    assert np.array_equal(nodes[:n - 1, 3], ref[:n - 1, 3])           # data[1] of the internal nodes


def test_sort_and_scan_64m(cuda_env):
    """configs[2]: 2^26 uint32 keys, keys-only and (key, index) pairs, and the scan of 2^26 values."""
    import torch
    from collision_b200 import _lib
    dev, cq = cuda_env
    L, P, S = _lib.lib(), _lib.ptr, cq.cuda_stream
    n = 1 << 26
    g = torch.Generator(device=dev).manual_seed(26)
    wide = torch.randint(0, 1 << 32, (n,), dtype=torch.int64, device=dev, generator=g)
    wide[: n // 8] &= 0xFFFF                 # a band of heavily duplicated keys: stability matters there
    keys = wide.to(torch.int32)              # the low 32 bits, as a bit pattern
    want_keys, want_idx = torch.sort(wide, stable=True)
    cq.wait_stream(torch.cuda.current_stream(dev))
    for value_bytes in (0, 4):
        scratch = _lib.alloc_bytes(dev, L.col_sort_scratch_bytes(n, 4, value_bytes))
        kout = torch.empty_like(keys)
        vin = torch.arange(n, dtype=torch.int32, device=dev) if value_bytes else None
        vout = torch.empty_like(keys) if value_bytes else None
        cq.wait_stream(torch.cuda.current_stream(dev))
        assert L.col_sort_pairs(P(keys), P(kout), P(vin), P(vout), n, 4, value_bytes, 0, P(scratch), S) == 0
        cq.synchronize()
        assert torch.equal(kout.to(torch.int64) & 0xFFFFFFFF, want_keys)     # sorted, and a permutation
        if value_bytes:
            assert torch.equal(vout.to(torch.int64), want_idx)               # stable
        assert torch.equal(keys.to(torch.int64) & 0xFFFFFFFF, wide)          # the input is left alone
        del scratch, kout, vin, vout
    del want_keys, want_idx
    data = (wide & 0xFF).to(torch.int32)
    want = torch.cumsum(data.to(torch.int64), 0)
    scratch = _lib.alloc_bytes(dev, L.col_scan_scratch_bytes(n))
    tot = torch.zeros(1, dtype=torch.int64, device=dev)
    cq.wait_stream(torch.cuda.current_stream(dev))
    assert L.col_exclusive_scan_u32(P(data), n, P(tot), P(scratch), S) == 0
    cq.synchronize()
    assert int(tot.item()) == int(want[-1].item())
    assert data[0].item() == 0
    assert torch.equal(data[1:].to(torch.int64) & 0xFFFFFFFF, want[:-1] & 0xFFFFFFFF)


def test_pipeline_128m_properties(cuda_env):
    """configs[4] on one GPU: 2^27 uniform spheres (the oracle would need minutes and tens of GB)."""
    import torch
    from collision_b200 import _lib
    dev, cq = cuda_env
    n = 1 << 27
    g = torch.Generator(device=dev).manual_seed(27)
    coords = torch.zeros(n, 4, device=dev)
    coords[:, :3] = torch.rand(n, 3, device=dev, generator=g)
    radii = torch.full((n,), 0.5 * n ** (-1.0 / 3.0), device=dev)
    from collision_b200.collision import Collider
    col = Collider(dev, n, 8, 128)
    n_buf = torch.zeros(1, dtype=torch.int32, device=dev)
    cq.wait_stream(torch.cuda.current_stream(dev))
    col.get_collisions(cq, coords, radii, n_buf, None, 0).synchronize()
    total = int(col._total_buf.item())
    assert 0 < total < 1 << 32 and (int(n_buf.item()) & 0xFFFFFFFF) == total
    pairs = torch.empty(total * 2, dtype=torch.int32, device=dev)
    col.get_collisions(cq, coords, radii, n_buf, pairs, total).synchronize()
    assert int(col._total_buf.item()) == total
    rows = pairs.view(-1, 2)
    # an independent count: the reference-layout kernel on the Node / Bound arrays
    L, P, S = _lib.lib(), _lib.ptr, cq.cuda_stream
    counter = torch.zeros(1, dtype=torch.int32, device=dev)
    assert L.col_traverse_atomic(None, P(counter), 0, P(col._nodes_buf), P(col._bounds_buf), 0, n, 0, n, S) == 0
    cq.synchronize()
    assert (int(counter.item()) & 0xFFFFFFFF) == total
    # sorted order: codes non-decreasing, ids a permutation
    codes = col._codes_bufs[1][:4 * n].view(torch.int32)
    assert bool((codes[1:] >= codes[:-1]).all())          # 30-bit codes: signed compare is fine
    ids = col._ids_bufs[1][:4 * n].view(torch.int32).to(torch.int64)
    where = torch.empty(n, dtype=torch.int32, device=dev)  # id -> sorted position
    where[ids] = torch.arange(n, dtype=torch.int32, device=dev)
    assert bool((ids[where.to(torch.int64)] == torch.arange(n, device=dev)).all())
    del codes
    # the list: grouped by query in sorted order, partner strictly later, boxes overlap (strictly, as collision.cl:164-166)
    step = 1 << 26
    last = -1
    for lo in range(0, total, step):
        a = rows[lo:lo + step, 0].to(torch.int64)
        b = rows[lo:lo + step, 1].to(torch.int64)
        pa, pb = where[a], where[b]
        assert bool((pb > pa).all())
        assert bool((pa[1:] >= pa[:-1]).all()) and int(pa[0].item()) >= last
        last = int(pa[-1].item())
        ca, cb = coords[a, :3], coords[b, :3]
        ra, rb = radii[a].unsqueeze(1), radii[b].unsqueeze(1)
        assert bool((((ca + ra) > (cb - rb)) & ((ca - ra) < (cb + rb))).all())
        del a, b, pa, pb, ca, cb, ra, rb
