"""Pins the CPU oracle to every known-answer vector the reference's own tests
hold for the hot path.  CPU only."""
import numpy as np
import pytest

from golden_data import *  # noqa: F401,F403
import golden_data as G


@pytest.mark.parametrize("dt", ["float32", "float64"])
def test_codes(oracle, dt):
    # tests/test_collision.py:251-298
    c = G.slots(G.SIX_COORDS, dt)
    rng = np.zeros((2, 4), dt)
    rng[0, :3] = G.SIX_COORDS.min(axis=0)
    rng[1, :3] = G.SIX_COORDS.max(axis=0)
    np.testing.assert_equal(oracle.morton(c, rng), G.SIX_CODES)


def test_fill_leaves(oracle):
    # tests/test_collision.py:50-75
    n = 8
    ids = np.random.RandomState(4).permutation(n).astype('uint32')
    nodes = np.zeros(2 * n - 1, G.Node)
    oracle.fill_leaves(nodes, ids)
    np.testing.assert_equal(nodes['data'][n - 1:, 0], ids)
    np.testing.assert_equal(nodes['right_edge'][n - 1:], np.arange(n))


@pytest.mark.parametrize("codes,expected", [(G.FIG3_CODES, G.fig3_expected), (G.ODD_CODES, G.odd_expected)])
def test_generate_bvh(oracle, codes, expected):
    # tests/test_collision.py:78-128, :131-179
    n = len(codes)
    nodes = oracle.build_tree(codes, np.arange(n, dtype='uint32'), fill=0xFFFFFFFF)
    exp, leaf_parents = expected()
    leaf = n - 1
    np.testing.assert_equal(nodes[1:leaf], exp[1:])
    np.testing.assert_equal(nodes[['data', 'right_edge']][0], exp[['data', 'right_edge']][0])
    np.testing.assert_equal(nodes['parent'][leaf:], leaf_parents)
    np.testing.assert_equal(nodes['right_edge'][leaf:], np.arange(n))
    np.testing.assert_equal(nodes['data'][leaf:, 0], np.arange(n))


def test_problem_codes(oracle):
    # tests/test_collision.py:425-480: every internal node but the root has a parent
    n = len(G.PROBLEM_CODES)
    nodes = oracle.build_tree(G.PROBLEM_CODES, np.arange(n, dtype='uint32'), fill=0xFFFFFFFF)
    parents = nodes['parent'][1:]
    assert (parents < n - 1).all()
    children = np.sort(nodes['data'][:n - 1].reshape(-1))
    np.testing.assert_equal(children, np.arange(1, 2 * n - 1))


@pytest.mark.parametrize("dt", ["float32", "float64"])
def test_compute_bounds(oracle, dt):
    # tests/test_collision.py:182-248
    c = G.slots(G.REFIT_COORDS, dt)
    r = np.ones(4, dt)
    b = oracle.refit(c, r, G.refit_nodes())
    np.testing.assert_equal(b[:, :, :3], G.REFIT_EXPECTED.astype(dt))


@pytest.mark.parametrize("dt", ["float32", "float64"])
def test_traverse(oracle, dt):
    # tests/test_collision.py:302-422 (host mergesort in place of the radix sort)
    c = G.slots(G.SIX_COORDS, dt)
    r = np.ones(6, dt)
    rng = np.zeros((2, 4), dt)
    rng[0, :3] = G.SIX_COORDS.min(axis=0)
    rng[1, :3] = G.SIX_COORDS.max(axis=0)
    codes = oracle.morton(c, rng)
    order = np.argsort(codes, kind='mergesort').astype('uint32')
    nodes = oracle.build_tree(codes[order], order)
    b = oracle.refit(c, r, nodes)
    total, pairs, _ = oracle.traverse(nodes, b, capacity=2)
    assert total == 2
    assert G.canon(pairs) == G.SIX_PAIRS


def test_collision_pipeline_six(oracle):
    # tests/test_collision_py.py:49-97 (N=6, ngroups=3, G=8 -> dropped-partial quirk harmless)
    c = G.slots(G.SIX_COORDS, 'float32')
    res = oracle.collide(c, np.ones(6, 'f4'), quirk_ngroups=(3, 8))
    assert res['total'] == 2 and G.canon(res['pairs']) == G.SIX_PAIRS
    res = oracle.collide(c, np.ones(6, 'f4'))
    assert res['total'] == 2 and G.canon(res['pairs']) == G.SIX_PAIRS


@pytest.mark.parametrize("size,ngroups,group_size", [(120, 5, 8), (256, 4, 32), (317, 4, 16), (341, 4, 64)])
def test_random_collision(oracle, size, ngroups, group_size):
    # tests/test_collision_py.py:100-150: oracle == brute-force box test, as row-sorted sets
    c, r = G.random_spheres(size, pad=np.nan)
    bf = oracle.brute_force(c, r)
    for quirk in (None, (ngroups, group_size)):
        res = oracle.collide(c, r, quirk_ngroups=quirk)
        assert res['total'] == len(bf)
        assert G.canon(res['pairs']) == G.canon(bf)
        # orientation: the second leaf always lies to the right in sorted order
        pos = np.empty(size, np.int64)
        pos[res['sorted_ids']] = np.arange(size)
        assert (pos[res['pairs'][:, 1]] > pos[res['pairs'][:, 0]]).all()


def test_pipeline_mt_matches_sequential(oracle):
    c, r = G.random_spheres(5000, pad=0.0)
    res = oracle.collide(c, r)
    p = oracle.Pipeline(5000)
    total, pairs = p.run(c, r, res['total'])
    assert total == res['total']
    np.testing.assert_equal(pairs, res['pairs'])
    np.testing.assert_equal(p.sorted_codes, res['sorted_codes'])
    np.testing.assert_equal(p.sorted_ids, res['sorted_ids'])
    np.testing.assert_equal(p.nodes[1:]['parent'], res['nodes'][1:]['parent'])
    np.testing.assert_equal(p.bounds[:, :, :3], res['bounds'][:, :, :3])
    total2, pairs2 = p.run(c, r, 10)   # capacity clamp keeps the first rows, total stays true
    assert total2 == total and np.array_equal(pairs2, res['pairs'][:10])


def test_bounds_quirk(oracle):
    # non-power-of-two ngroups drops tail partials (reduce.cl:42-52)
    v = np.zeros((10, 4), 'f4')
    v[:, 0] = np.arange(10)
    exact = oracle.bounds(v)
    assert exact[0, 0] == 0 and exact[1, 0] == 9
    q = oracle.bounds(v, ngroups=3, group_size=1)      # threads 0..2 stride 3; partial 2 dropped
    assert q[1, 0] == 9 and q[0, 0] == 0
    v[:, 0] = [0, 1, 50, 3, 4, 5, 6, 7, 8, 9]          # max lives only in group 2's partial (i = 2, 5, 8)
    q = oracle.bounds(v, ngroups=3, group_size=1)
    assert q[1, 0] == 9.0 and oracle.bounds(v)[1, 0] == 50.0
    np.testing.assert_equal(oracle.bounds(v, ngroups=4, group_size=2), oracle.bounds(v))


@pytest.mark.parametrize("dt,lanes", [("float32", 4), ("float64", 4), ("float32", 1)])
def test_bounds_random(oracle, dt, lanes):
    # tests/test_bounds_py.py:18-148
    rs = np.random.RandomState(4)
    for scale, shift in ((1.0, 0.0), (1e8, 0.0), (1.0, -2.0)):
        v = (rs.random_sample((1000, lanes)) * scale + shift).astype(dt)
        out = oracle.bounds(v, 8, 16)
        np.testing.assert_equal(out[0], v.min(axis=0))
        np.testing.assert_equal(out[1], v.max(axis=0))


def test_scan_golden(oracle):
    # tests/test_scan.py:24-104 (two 8-blocks -> block sums [118,119] -> full scan)
    values = np.array([17, 6, 24, 28, 18, 22, 2, 1, 25, 17, 7, 17, 3, 19, 8, 23], dtype='uint32')
    full = np.array([0, 17, 23, 47, 75, 93, 115, 117, 118, 143, 160, 167, 184, 187, 206, 214], dtype='uint32')
    np.testing.assert_equal(oracle.exclusive_scan(values), full)
    np.testing.assert_equal(oracle.exclusive_scan(values[:8]), full[:8])
    assert values[:8].sum() == 118 and values[8:].sum() == 119
    rs = np.random.RandomState(4)
    for size in (20, 24, 1024, 160, 320):       # tests/test_scan_py.py:47-64
        v = rs.randint(0, size, size=size).astype('uint32')
        exp = np.cumsum(v)
        out = oracle.exclusive_scan(v)
        assert out[0] == 0
        np.testing.assert_equal(out[1:], exp[:-1])
    big = np.full(8, 0xC0000000, 'uint32')      # wraps mod 2^32
    np.testing.assert_equal(oracle.exclusive_scan(big), (np.arange(8, dtype=np.uint64) * 0xC0000000 % 2**32).astype('u4'))


@pytest.mark.parametrize("kdt", ["uint32", "uint64"])
@pytest.mark.parametrize("vdt", [None, "uint32", "float64", ("float64", 4)])
def test_sort(oracle, kdt, vdt):
    # tests/test_radix.py:185-351, tests/test_radix_py.py:83-201: == stable argsort, every pass too
    rs = np.random.RandomState(4)
    n = 15360
    keys = rs.randint(0, 64, size=n).astype(kdt)
    keys[::7] = rs.randint(0, 2**31, size=len(keys[::7]))
    values = None if vdt is None else (rs.random_sample((n,) + ((4,) if isinstance(vdt, tuple) else ())) * 1000).astype(
        vdt[0] if isinstance(vdt, tuple) else vdt)
    order = np.argsort(keys, kind='stable')
    ko, vo = oracle.sort_pairs(keys, values)
    np.testing.assert_equal(ko, keys[order])
    if values is not None:
        np.testing.assert_equal(vo, values[order])
    k, v = keys, values
    for p in range(keys.dtype.itemsize * 8 // 4):      # 4-bit passes as the reference does
        digit = (k >> np.array(4 * p, k.dtype)) & np.array(15, k.dtype)
        o = np.argsort(digit, kind='stable')
        k2, v2 = oracle.radix_pass(k, v, 4, p)
        np.testing.assert_equal(k2, k[o])
        if v is not None:
            np.testing.assert_equal(v2, v[o])
        k, v = k2, v2
    np.testing.assert_equal(k, keys[order])


def test_offsets(oracle):
    # tests/test_offset.py:31-47, tests/test_offset_py.py:23-62
    np.testing.assert_equal(oracle.find_offsets(G.OFFSET_VALUES, 7), G.OFFSET_EXPECTED)
    missing = np.array([1, 1, 1, 1, 1, 1, 1, 3, 3, 3, 3, 3, 3], dtype='uint32')
    np.testing.assert_equal(oracle.find_offsets(missing, 7), [0, 0, 7, 7, 13, 13, 13])
    rs = np.random.RandomState(4)
    v = np.sort(rs.randint(0, 500, size=4096)).astype('uint64')
    out = oracle.find_offsets(v, 600, np.uint64)
    np.testing.assert_equal(out, np.searchsorted(v, np.arange(600), side='left'))


def test_gather_scatter(oracle):
    # tests/test_index.py:18-79
    rs = np.random.RandomState(4)
    vals = rs.random_sample(100)
    idx = rs.permutation(100)[:60].astype('uint64')
    out = np.full(60, -1.0)
    np.testing.assert_equal(oracle.gather(vals, idx, out), vals[idx])
    out = np.full(100, -1.0)
    oracle.scatter(vals[:60].copy(), idx, out)
    exp = np.full(100, -1.0)
    exp[idx] = vals[:60]
    np.testing.assert_equal(out, exp)
