"""GPU parity tests of the LBVH path: every stage and the whole pipeline against
the CPU oracle and the reference's golden vectors.  Mirrors upstream's
tests/test_collision.py (kernel level) and tests/test_collision_py.py (object level).
All calls go through the C ABI (via collision_b200's ctypes binding)."""
import numpy as np
import pytest

import golden_data as G
from gpu_util import dev_empty, to_dev, to_host

pytestmark = pytest.mark.gpu

Node = G.Node


@pytest.fixture(scope="module", params=["float32", "float64"])
def kernels(request, cuda_env):
    from collision_b200.collision import CollisionProgram
    dev, cq = cuda_env
    return CollisionProgram(dev, request.param).kernels, np.dtype(request.param)


def scene_range(coords, dt):
    rng = np.zeros((2, 4), dt)
    rng[0, :3] = coords.min(axis=0)
    rng[1, :3] = coords.max(axis=0)
    return rng


def test_fill_internal(cuda_env, kernels):
    # tests/test_collision.py:50-75
    dev, cq = cuda_env
    k, _ = kernels
    n = 8
    ids = np.random.RandomState(4).permutation(n).astype('uint32')
    nodes_buf = dev_empty((2 * n - 1) * 16, fill=0)
    k['fillInternal'](cq, (32,), None, nodes_buf, to_dev(ids), n).synchronize()
    nodes = to_host(nodes_buf, Node, (2 * n - 1,))
    np.testing.assert_equal(nodes['data'][n - 1:, 0], ids)
    np.testing.assert_equal(nodes['right_edge'][n - 1:], np.arange(n))


@pytest.mark.parametrize("codes,expected", [(G.FIG3_CODES, G.fig3_expected), (G.ODD_CODES, G.odd_expected)])
def test_generate_bvh(cuda_env, kernels, codes, expected):
    # tests/test_collision.py:78-179
    dev, cq = cuda_env
    k, _ = kernels
    n = len(codes)
    nodes_buf = dev_empty((2 * n - 1) * 16, fill=0)
    e = k['fillInternal'](cq, (32,), None, nodes_buf, to_dev(np.arange(n, dtype='uint32')), n)
    k['generateBVH'](cq, (32,), None, to_dev(codes), nodes_buf, n, wait_for=[e]).synchronize()
    nodes = to_host(nodes_buf, Node, (2 * n - 1,))
    exp, leaf_parents = expected()
    leaf = n - 1
    np.testing.assert_equal(nodes[1:leaf], exp[1:])
    np.testing.assert_equal(nodes[['data', 'right_edge']][0], exp[['data', 'right_edge']][0])
    np.testing.assert_equal(nodes['parent'][leaf:], leaf_parents)
    np.testing.assert_equal(nodes['right_edge'][leaf:], np.arange(n))
    np.testing.assert_equal(nodes['data'][leaf:, 0], np.arange(n))


def test_problem_codes(cuda_env, kernels, oracle):
    # tests/test_collision.py:425-480
    dev, cq = cuda_env
    k, _ = kernels
    n = len(G.PROBLEM_CODES)
    ids = np.arange(n, dtype='uint32')
    nodes_buf = dev_empty((2 * n - 1) * 16, fill=0xFF)
    k['fillInternal'](cq, (32,), None, nodes_buf, to_dev(ids), n)
    k['generateBVH'](cq, (32,), None, to_dev(G.PROBLEM_CODES), nodes_buf, n).synchronize()
    nodes = to_host(nodes_buf, Node, (2 * n - 1,))
    want = oracle.build_tree(G.PROBLEM_CODES, ids, fill=0xFFFFFFFF)
    np.testing.assert_equal(nodes, want)
    assert (nodes['parent'][1:] < n - 1).all()


def test_compute_bounds(cuda_env, kernels):
    # tests/test_collision.py:182-248
    dev, cq = cuda_env
    k, dt = kernels
    c = G.slots(G.REFIT_COORDS, dt)
    nodes = G.refit_nodes()
    bounds_buf = dev_empty(len(nodes) * 8 * dt.itemsize, fill=0)
    flags_buf = dev_empty(len(nodes) * 4, fill=0xFF)   # the refit entry point clears them itself
    nodes_buf = to_dev(nodes)
    e = k['leafBounds'](cq, (32,), None, bounds_buf, to_dev(c), to_dev(np.ones(4, dt)), nodes_buf, 4)
    k['internalBounds'](cq, (32,), None, bounds_buf, flags_buf, nodes_buf, 4, wait_for=[e]).synchronize()
    b = to_host(bounds_buf, dt, (len(nodes), 2, 4))
    np.testing.assert_equal(b[:, :, :3], G.REFIT_EXPECTED.astype(dt))


def test_codes(cuda_env, kernels):
    # tests/test_collision.py:251-298
    dev, cq = cuda_env
    k, dt = kernels
    c = G.slots(G.SIX_COORDS, dt)
    codes_buf = dev_empty(6 * 4)
    k['calculateCodes'](cq, (32,), None, codes_buf, to_dev(c), to_dev(scene_range(G.SIX_COORDS, dt)), 6).synchronize()
    np.testing.assert_equal(to_host(codes_buf, 'uint32', (6,)), G.SIX_CODES)


def test_traverse(cuda_env, kernels):
    # tests/test_collision.py:302-422: the reference's own kernel sequence, host mergesort for the sort
    dev, cq = cuda_env
    k, dt = kernels
    n = 6
    c = G.slots(G.SIX_COORDS, dt)
    coords_buf, radii_buf = to_dev(c), to_dev(np.ones(n, dt))
    codes_buf = dev_empty(n * 4)
    k['calculateCodes'](cq, (32,), None, codes_buf, coords_buf, to_dev(scene_range(G.SIX_COORDS, dt)), n).synchronize()
    codes = to_host(codes_buf, 'uint32', (n,))
    order = np.argsort(codes, kind='mergesort').astype('uint32')
    codes_buf = to_dev(codes[order])
    nodes_buf = dev_empty((2 * n - 1) * 16, fill=0)
    bounds_buf = dev_empty((2 * n - 1) * 8 * dt.itemsize, fill=0)
    flags_buf = dev_empty((2 * n - 1) * 4, fill=0)
    collisions_buf = dev_empty(2 * 2 * 4, fill=0xFF)
    n_buf = dev_empty(4, fill=0)
    k['fillInternal'](cq, (32,), None, nodes_buf, to_dev(order), n)
    k['generateBVH'](cq, (32,), None, codes_buf, nodes_buf, n)
    k['leafBounds'](cq, (32,), None, bounds_buf, coords_buf, radii_buf, nodes_buf, n)
    k['internalBounds'](cq, (32,), None, bounds_buf, flags_buf, nodes_buf, n)
    k['traverse'](cq, (32,), None, collisions_buf, n_buf, 2, nodes_buf, bounds_buf, n).synchronize()
    assert to_host(n_buf, 'uint32', (1,))[0] == 2
    assert G.canon(to_host(collisions_buf, 'uint32', (2, 2))) == G.SIX_PAIRS


# ------------------------------------------------------------ object level
def run_collider(cuda_env, collider, c, r, capacity, dt):
    dev, cq = cuda_env
    coords_buf, radii_buf = to_dev(c), to_dev(r.astype(dt))
    n_buf = dev_empty(4, fill=0xFF)
    col_buf = dev_empty(max(capacity, 1) * 8, fill=0xFF) if capacity else None
    collider.get_collisions(cq, coords_buf, radii_buf, n_buf, col_buf, capacity).synchronize()
    total = int(to_host(n_buf, 'uint32', (1,))[0])
    pairs = to_host(col_buf, 'uint32', (capacity, 2))[:min(total, capacity)] if capacity else np.zeros((0, 2), 'u4')
    return total, pairs


def test_padded_size(cuda_env):
    # tests/test_collision_py.py:40-46
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    for size, ngroups, group_size, expected in [(48, 3, 8, 48), (47, 3, 8, 48), (49, 3, 8, 64)]:
        assert Collider(dev, size, ngroups, group_size).padded_size == expected


@pytest.mark.parametrize("dt", ["float32", "float64"])
@pytest.mark.parametrize("ordered", [True, False])
def test_collision(cuda_env, dt, ordered):
    # tests/test_collision_py.py:49-97
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    collider = Collider(dev, 6, 3, 8, np.dtype(dt), ordered=ordered)
    total, pairs = run_collider(cuda_env, collider, G.slots(G.SIX_COORDS, dt), np.ones(6), 2, np.dtype(dt))
    assert total == 2 and G.canon(pairs) == G.SIX_PAIRS


@pytest.mark.parametrize("size,ngroups,group_size", [(120, 5, 8), (256, 4, 32), (317, 4, 16), (341, 4, 64),
                                                     (2, 1, 8), (3, 1, 8), (5000, 8, 64)])
@pytest.mark.parametrize("dt", ["float32", "float64"])
@pytest.mark.parametrize("layout", ["wide", "reference"])
def test_random_collision(cuda_env, oracle, size, ngroups, group_size, dt, layout):
    # tests/test_collision_py.py:100-150 -- plus stage-by-stage bit parity with the oracle
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    dt = np.dtype(dt)
    c, r = G.random_spheres(size, dtype=dt, pad=np.nan)
    want = oracle.collide(c, r)
    collider = Collider(dev, size, ngroups, group_size, dt, layout=layout)
    cap = want['total'] + 3
    total, pairs = run_collider(cuda_env, collider, c, r, cap, dt)
    n = size
    np.testing.assert_equal(to_host(collider._codes_bufs[1], 'uint32', (n,)), want['sorted_codes'])
    np.testing.assert_equal(to_host(collider._ids_bufs[1], 'uint32', (n,)), want['sorted_ids'])
    nodes = to_host(collider._nodes_buf, Node, (2 * n - 1,))
    np.testing.assert_equal(nodes[1:]['parent'], want['nodes'][1:]['parent'])
    np.testing.assert_equal(nodes['right_edge'], want['nodes']['right_edge'])
    np.testing.assert_equal(nodes['data'][:n - 1], want['nodes']['data'][:n - 1])
    np.testing.assert_equal(nodes['data'][n - 1:, 0], want['nodes']['data'][n - 1:, 0])
    b = to_host(collider._bounds_buf, dt, (2 * n - 1, 2, 4))
    np.testing.assert_equal(b[:, :, :3], want['bounds'][:, :, :3])
    assert total == want['total']
    np.testing.assert_equal(pairs, want['pairs'])          # identical ORDER (deterministic compaction)
    assert G.canon(pairs) == G.canon(oracle.brute_force(c, r))
    # reference-style single pass: same set, arbitrary order
    total2, pairs2 = run_collider(cuda_env, Collider(dev, size, ngroups, group_size, dt, ordered=False), c, r, cap, dt)
    assert total2 == total
    np.testing.assert_equal(G.lexsorted(pairs2), G.lexsorted(want['pairs']))


@pytest.mark.parametrize("old,new", [((350, 8, 64), (351, 8, 64)), ((350, 8, 64), (351, None, None)),
                                     ((351, 8, 64), (100, 4, 16))])
def test_random_collision_resized(cuda_env, oracle, old, new):
    # tests/test_collision_py.py:153-207
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    collider = Collider(dev, *old)
    collider.resize(*new)
    size = new[0]
    c, r = G.random_spheres(size, pad=np.nan)
    want = oracle.collide(c, r)
    total, pairs = run_collider(cuda_env, collider, c, r, want['total'], np.dtype('float32'))
    assert total == want['total']
    np.testing.assert_equal(pairs, want['pairs'])


@pytest.mark.parametrize("ordered,layout", [(True, "wide"), (True, "reference"), (False, "reference")])
def test_count_only_and_overflow(cuda_env, oracle, ordered, layout):
    # tests/test_collision_py.py:261-327
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    size = 1000
    c, r = G.random_spheres(size)
    want = oracle.collide(c, r)
    collider = Collider(dev, size, 10, 32, ordered=ordered, layout=layout)
    total, _ = run_collider(cuda_env, collider, c, r, 0, np.dtype('float32'))
    assert total == want['total']
    with pytest.raises(ValueError):
        collider.get_collisions(cq, to_dev(c), to_dev(r), dev_empty(4), None, 5)
    # capacity smaller than the total: counter still the true total, buffer not overrun
    cap = 7
    col_buf = dev_empty((cap + 4) * 8, fill=0xFF)
    n_buf = dev_empty(4)
    collider.get_collisions(cq, to_dev(c), to_dev(r), n_buf, col_buf, cap).synchronize()
    assert int(to_host(n_buf, 'uint32', (1,))[0]) == want['total']
    got = to_host(col_buf, 'uint32', (cap + 4, 2))
    assert (got[cap:] == 0xFFFFFFFF).all()
    if ordered:
        np.testing.assert_equal(got[:cap], want['pairs'][:cap])
    else:
        assert G.canon(got[:cap]) <= G.canon(want['pairs'])


@pytest.mark.parametrize("shape", G.AWKWARD_SHAPES)
def test_awkward_scenes(cuda_env, oracle, shape):
    """Duplicate Morton codes (repeated points, lattices), collinear and coplanar centres, radii from zero
    to scene-sized -- the scenes on which the oracle is pinned to the reference's own kernel text
    (tests/test_oracle_vs_ref.py): every stage and the ordered rows bit for bit."""
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    dt = np.dtype('float32')
    for seed in range(4):
        c, r = G.awkward_scene(shape, seed, dt)
        n = len(c)
        want = oracle.collide(c, r)
        collider = Collider(dev, n, 4, 16, dt)
        total, pairs = run_collider(cuda_env, collider, c, r, want['total'] + 3, dt)
        np.testing.assert_equal(to_host(collider._codes_bufs[1], 'uint32', (n,)), want['sorted_codes'])
        np.testing.assert_equal(to_host(collider._ids_bufs[1], 'uint32', (n,)), want['sorted_ids'])
        nodes = to_host(collider._nodes_buf, Node, (2 * n - 1,))
        np.testing.assert_equal(nodes[1:]['parent'], want['nodes'][1:]['parent'])
        np.testing.assert_equal(nodes['data'][:n - 1], want['nodes']['data'][:n - 1])
        b = to_host(collider._bounds_buf, dt, (2 * n - 1, 2, 4))
        np.testing.assert_equal(b[:, :, :3], want['bounds'][:, :, :3])
        assert total == want['total']
        np.testing.assert_equal(pairs, want['pairs'])


def test_program_mismatch(cuda_env):
    # tests/test_collision_py.py:329-336 style: shared programs must match
    from collision_b200.collision import Collider, CollisionProgram
    dev, cq = cuda_env
    prog = CollisionProgram(dev, 'float64')
    with pytest.raises(ValueError):
        Collider(dev, 64, 4, 8, np.dtype('float32'), program=prog)
    c = Collider(dev, 64, 4, 8, np.dtype('float64'), program=prog)
    assert c.program is prog
    with pytest.raises(ValueError):
        CollisionProgram(dev, 'int32')


@pytest.mark.parametrize("dt", ["float32", "float64"])
def test_dense_cluster_overflows_stash(cuda_env, oracle, dt):
    """Many more than 32 hits per query (the fused kernel's stash) and heavy code
    duplication: exercises the second walk and the duplicate tie-break."""
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    dt = np.dtype(dt)
    rs = np.random.RandomState(7)
    n = 3000
    c = np.zeros((n, 4), dt)
    c[:, :3] = rs.normal(0.5, 0.01, (n, 3))
    c[::3] = c[1::3][:len(c[::3])]                      # exact duplicates
    r = np.full(n, 0.004, dt)
    want = oracle.collide(c, r)
    assert want['counts'].max() > 64
    total, pairs = run_collider(cuda_env, Collider(dev, n, 8, 64, dt), c, r, want['total'], dt)
    assert total == want['total']
    np.testing.assert_equal(pairs, want['pairs'])
    # a range-limited capacity still yields the first rows
    total, pairs = run_collider(cuda_env, Collider(dev, n, 8, 64, dt), c, r, 1000, dt)
    assert total == want['total']
    np.testing.assert_equal(pairs, want['pairs'][:1000])


def test_wide_count_and_write_entry_points(cuda_env, oracle):
    """col_traverse_wide_count / _write over query sub-ranges (the multi-GPU sharding primitives)."""
    import torch
    from collision_b200 import _lib
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    n = 4000
    c, r = G.random_spheres(n)
    want = oracle.collide(c, r)
    col = Collider(dev, n, 8, 64)
    run_collider(cuda_env, col, c, r, 0, np.dtype('float32'))
    L, P, S = _lib.lib(), _lib.ptr, cq.cuda_stream
    bounds = [0, 1000, 1001, 2500, n]
    rows = []
    for qb, qe in zip(bounds, bounds[1:]):
        counts = dev_empty((qe - qb) * 4)
        tot = torch.zeros(1, dtype=torch.int64, device=dev)
        assert L.col_traverse_wide_count(P(tot), P(counts), P(col._wide_buf), P(col._bounds_buf), P(col._scene_buf), 0, n, qb, qe, S) == 0
        cq.synchronize()
        cnt = to_host(counts, 'uint32', (qe - qb,))
        np.testing.assert_equal(cnt, want['counts'][qb:qe])
        assert int(tot.item()) == int(cnt.sum())
        offs = to_dev(oracle.exclusive_scan(cnt))
        out = dev_empty(max(int(cnt.sum()), 1) * 8)
        assert L.col_traverse_wide_write(P(out), int(cnt.sum()), P(offs), 0, P(col._wide_buf), P(col._bounds_buf),
                                         P(col._scene_buf), P(col._ids_bufs[1]), 0, n, qb, qe, S) == 0
        cq.synchronize()
        rows.append(to_host(out, 'uint32', (int(cnt.sum()), 2)))
        # fused over the same sub-range gives the same rows
        out2 = dev_empty(max(int(cnt.sum()), 1) * 8)
        scr = dev_empty(L.col_traverse_scratch_bytes(qe - qb, int(cnt.sum())))
        assert L.col_traverse_wide_pairs(P(out2), int(cnt.sum()), 0, P(tot), P(col._wide_buf), P(col._bounds_buf),
                                         P(col._scene_buf), P(col._ids_bufs[1]), 0, n, qb, qe, P(scr), S) == 0
        cq.synchronize()
        np.testing.assert_equal(to_host(out2, 'uint32', (int(cnt.sum()), 2)), rows[-1])
        assert int(tot.item()) == int(cnt.sum())
    np.testing.assert_equal(np.concatenate(rows), want['pairs'])


def test_graph_replay_matches_plain_enqueue(cuda_env, oracle):
    """Collider(graph=True): the captured launch sequence replays to the same rows, for
    changing inputs in the same buffers, and is re-captured when the capacity changes."""
    import numpy as np
    import torch
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    n = 5000
    col = Collider(dev, n, 8, 128, graph=True)
    n_buf = torch.zeros(1, dtype=torch.int32, device=dev)
    coords_d = torch.zeros(n, 4, dtype=torch.float32, device=dev)
    radii_d = torch.zeros(n, dtype=torch.float32, device=dev)
    cap = 60000
    pairs_d = torch.zeros(cap * 2, dtype=torch.int32, device=dev)
    for seed in range(5):
        rs = np.random.RandomState(seed)
        coords = np.zeros((n, 4), np.float32)
        coords[:, :3] = rs.random_sample((n, 3))
        radii = np.full(n, 0.02 + 0.002 * seed, np.float32)
        want = oracle.collide(coords, radii)
        coords_d.copy_(torch.from_numpy(coords))
        radii_d.copy_(torch.from_numpy(radii))
        torch.cuda.synchronize()
        col.get_collisions(cq, coords_d, radii_d, n_buf, pairs_d, cap).synchronize()
        total = int(n_buf.item()) & 0xFFFFFFFF
        assert total == want["total"] and total <= cap
        got = pairs_d.cpu().numpy().view(np.uint32).reshape(-1, 2)[:total]
        assert np.array_equal(got, want["pairs"])
    assert len(col._graphs) == 1
    # a different capacity is a different launch sequence
    col.get_collisions(cq, coords_d, radii_d, n_buf, pairs_d, cap // 2).synchronize()
    col.get_collisions(cq, coords_d, radii_d, n_buf, pairs_d, cap // 2).synchronize()
    assert (int(n_buf.item()) & 0xFFFFFFFF) == want["total"]


@pytest.mark.parametrize("dt", ["float32", "float64"])
@pytest.mark.parametrize("n,dup", [(1, False), (2, False), (3, True), (257, False), (1000, True), (70001, False),
                                   (70001, True), ((1 << 22) + 5, False)])   # (the last: two-pass form of the fused build)
def test_fused_build_matches_tree_plus_refit(cuda_env, oracle, n, dup, dt):
    """col_build_refit_wide (tree found bottom-up inside the refit climb) writes byte-identical
    Node, Bound and traversal-record arrays to col_build_tree_wide + col_refit_wide (Karras'
    searches, collision.cl:72-121), and the Node array equals the oracle's -- including runs
    of equal Morton codes, where the tie-break on the sorted position decides the topology."""
    import torch
    from collision_b200 import _lib
    dev, cq = cuda_env
    dt = np.dtype(dt)
    code = _lib.float_code(dt)
    L, P, S = _lib.lib(), _lib.ptr, cq.cuda_stream
    rs = np.random.RandomState(n + dup)
    coords = np.zeros((n, 4), dt)
    coords[:, :3] = rs.random_sample((n, 3))
    radii = (rs.random_sample(n) * 0.02).astype(dt)
    codes = rs.randint(0, 1 << 30, n).astype(np.uint32)
    if dup:  # long runs of identical codes
        codes = (codes >> 22) << 22
        codes[: n // 3] = codes[0]
    codes.sort()
    ids = rs.permutation(n).astype(np.uint32)
    scene = np.zeros((2, 4), dt)
    scene[1, :3] = 1
    d_codes, d_ids, d_coords, d_radii, d_scene = map(to_dev, (codes, ids, coords, radii, scene))
    got = []
    for fused in (False, True):
        nodes = torch.full(((2 * n - 1) * 16,), 0xAB, dtype=torch.uint8, device=dev)
        bounds = torch.zeros((2 * n - 1) * 8 * dt.itemsize, dtype=torch.uint8, device=dev)
        wide = torch.zeros(max(n - 1, 1) * 32, dtype=torch.uint8, device=dev)
        flags = torch.full((2 * n * 4,), 0x5A, dtype=torch.uint8, device=dev)   # (the fused call wants 2n words)
        cq.wait_stream(torch.cuda.current_stream(dev))   # the fills above run on torch's current stream
        if fused:
            assert L.col_build_refit_wide(P(d_codes), P(d_ids), P(nodes), P(wide), P(flags), P(bounds), P(d_coords),
                                          P(d_radii), P(d_scene), code, n, S) == 0
        else:
            assert L.col_build_tree_wide(P(d_codes), P(d_ids), P(nodes), P(wide), P(flags), n, S) == 0
            assert L.col_refit_wide(P(bounds), P(wide), P(flags), P(nodes), P(d_coords), P(d_radii), P(d_scene),
                                    code, n, S) == 0
        cq.synchronize()
        got.append((nodes.cpu().numpy(), bounds.cpu().numpy(), wide.cpu().numpy()))
    for a, b, what in zip(got[0], got[1], ("nodes", "bounds", "records")):
        assert np.array_equal(a, b), what
    want = oracle.build_tree(codes, ids, fill=0xFFFFFFFF)   # untouched fields are NO_NODE in both
    np.testing.assert_equal(got[1][0].view(np.uint32).reshape(-1, 4), want.view(np.uint32).reshape(-1, 4))


def test_collider_two_launch_build(cuda_env, oracle):
    """Collider(fused_build=False) -- Karras' searches, then the climb -- gives the same rows."""
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    n = 6000
    c, r = G.random_spheres(n)
    want = oracle.collide(c, r)
    for fused in (True, False):
        col = Collider(dev, n, 8, 64, fused_build=fused)
        total, pairs = run_collider(cuda_env, col, c, r, want["total"], np.dtype("float32"))
        assert total == want["total"]
        np.testing.assert_equal(pairs, want["pairs"])


def test_configs0_exact(cuda_env, oracle):
    """BASELINE.json configs[0] exactly: np.random.seed(4), 1,024 spheres
    U[0,1)^3, r = 0.01, Collider(ctx, 1024, ngroups=8, group_size=64) -- every stage and the
    ordered list against the oracle (tests/test_collision_py.py:100-150 shape)."""
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    np.random.seed(4)
    n = 1024
    c = np.zeros((n, 4), np.float32)
    c[:, :3] = np.random.random((n, 3)).astype('float32')
    r = np.full(n, 0.01, np.float32)
    want = oracle.collide(c, r)
    assert 10 < want['total'] < 100          # ~ N^2/2 * (4r)^3 = 34
    collider = Collider(dev, n, 8, 64)
    assert collider.padded_size == 1024 and collider.n_nodes == 2047
    total, pairs = run_collider(cuda_env, collider, c, r, want['total'], np.dtype('float32'))
    np.testing.assert_equal(to_host(collider._codes_bufs[0], 'uint32', (n,)), want['codes'])
    np.testing.assert_equal(to_host(collider._codes_bufs[1], 'uint32', (n,)), want['sorted_codes'])
    np.testing.assert_equal(to_host(collider._ids_bufs[1], 'uint32', (n,)), want['sorted_ids'])
    nodes = to_host(collider._nodes_buf, Node, (2 * n - 1,))
    np.testing.assert_equal(nodes[1:]['parent'], want['nodes'][1:]['parent'])
    np.testing.assert_equal(nodes['right_edge'], want['nodes']['right_edge'])
    np.testing.assert_equal(nodes['data'][:n - 1], want['nodes']['data'][:n - 1])
    np.testing.assert_equal(nodes['data'][n - 1:, 0], want['nodes']['data'][n - 1:, 0])
    b = to_host(collider._bounds_buf, 'float32', (2 * n - 1, 2, 4))
    np.testing.assert_equal(b[:, :, :3], want['bounds'][:, :, :3])
    assert total == want['total']
    np.testing.assert_equal(pairs, want['pairs'])
    assert G.canon(pairs) == G.canon(oracle.brute_force(c, r))
    # count-only, and the reference's own unordered contract on the Node / Bound arrays
    total0, _ = run_collider(cuda_env, collider, c, r, 0, np.dtype('float32'))
    assert total0 == want['total']
    t2, p2 = run_collider(cuda_env, Collider(dev, n, 8, 64, ordered=False), c, r, want['total'], np.dtype('float32'))
    assert t2 == total
    np.testing.assert_equal(G.lexsorted(p2), G.lexsorted(want['pairs']))


def test_graph_recaptured_across_resize(cuda_env, oracle):
    """Collider(graph=True) across resize() to other sizes (re-capture on size
    change): every size replays its own capture and returns that size's rows."""
    import torch
    from collision_b200.collision import Collider
    dev, cq = cuda_env
    col = Collider(dev, 3000, 8, 64, graph=True)
    cap = 40000
    n_buf = torch.zeros(1, dtype=torch.int32, device=dev)
    pairs_d = torch.zeros(cap * 2, dtype=torch.int32, device=dev)
    for n in (3000, 3001, 777, 3000):
        col.resize(n)
        rs = np.random.RandomState(n)
        coords = np.zeros((n, 4), np.float32)
        coords[:, :3] = rs.random_sample((n, 3))
        radii = np.full(n, 0.03, np.float32)
        want = oracle.collide(coords, radii)
        coords_d, radii_d = torch.from_numpy(coords).to(dev), torch.from_numpy(radii).to(dev)
        for rep in range(3):      # plain, capture, replay
            pairs_d.zero_()
            torch.cuda.synchronize()
            col.get_collisions(cq, coords_d, radii_d, n_buf, pairs_d, cap).synchronize()
            total = int(n_buf.item()) & 0xFFFFFFFF
            assert total == want["total"] and total <= cap
            got = pairs_d.cpu().numpy().view(np.uint32).reshape(-1, 2)[:total]
            assert np.array_equal(got, want["pairs"]), (n, rep)
        assert len(col._graphs) == 1      # the captures of the previous size were dropped with its scratch
