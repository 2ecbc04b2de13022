"""Validates the C restatement against the reference's OWN OpenCL kernel text
(collision.cl / index.cl / offset.cl compiled as C++ by oracle/ref_shim.cpp into
oracle/_ref/).  Skipped where oracle/_ref was not built (no reference checkout)."""
import ctypes as C

import numpy as np
import pytest

import golden_data as G


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.fixture(params=["float32", "float64"])
def ref_env(request, oracle):
    lib = oracle.ref(request.param)
    if lib is None:
        pytest.skip("oracle/_ref not built")
    assert lib.ref_dtype_bytes() == np.dtype(request.param).itemsize
    return lib, request.param


def ref_pipeline(lib, dt, c, r, rng, capacity):
    """The reference kernels in Collider.get_collisions order (collision/collision.py:130-198),
    with numpy's stable argsort standing in for the radix sort (as tests/test_collision.py:368 does)."""
    n = len(c)
    gsz = C.c_size_t((n + 31) // 32 * 32)
    codes = np.zeros(n, np.uint32)
    lib.ref_calculateCodes(gsz, _p(codes), _p(c), _p(rng), C.c_uint(n))
    order = np.argsort(codes, kind='stable').astype(np.uint32)
    scodes = codes[order]
    nodes = np.zeros(2 * n - 1, G.Node)
    lib.ref_fillInternal(gsz, _p(nodes), _p(order), C.c_uint(n))
    lib.ref_generateBVH(C.c_size_t((n - 1 + 31) // 32 * 32), _p(scodes), _p(nodes), C.c_uint(n))
    b = np.zeros((2 * n - 1, 2, 4), dt)
    flags = np.zeros(2 * n - 1, np.uint32)
    lib.ref_leafBounds(gsz, _p(b), _p(c), _p(r), _p(nodes), C.c_uint(n))
    lib.ref_internalBounds(gsz, _p(b), _p(flags), _p(nodes), C.c_uint(n))
    pairs = np.zeros((max(capacity, 1), 2), np.uint32)
    nxt = np.zeros(1, np.uint32)
    lib.ref_traverse(gsz, _p(pairs), _p(nxt), C.c_uint(capacity), _p(nodes), _p(b), C.c_uint(n))
    return codes, order, nodes, b, int(nxt[0]), pairs[:min(capacity, int(nxt[0]))]


@pytest.mark.parametrize("n,clustered", [(2, False), (3, False), (257, False), (4096, False), (3000, True)])
def test_oracle_equals_reference_kernels(oracle, ref_env, n, clustered):
    lib, dt = ref_env
    rs = np.random.RandomState(4 + n)
    c = np.zeros((n, 4), dt)
    if clustered:   # duplicates + dense clusters: exercises the tie-break and deep stacks
        centres = rs.random_sample((8, 3))
        c[:, :3] = centres[rs.randint(0, 8, n)] + rs.normal(0, 0.01, (n, 3))
        c[::5, :3] = c[1::5, :3][:len(c[::5])]
    else:
        c[:, :3] = rs.random_sample((n, 3))
    r = rs.uniform(0, 1 / np.sqrt(n), n).astype(dt)
    compare_with_reference_kernels(oracle, lib, dt, c, r)


def compare_with_reference_kernels(oracle, lib, dt, c, r):
    rng = oracle.bounds(c)
    want = oracle.collide(c, r)
    codes, order, nodes, b, total, pairs = ref_pipeline(lib, dt, c, r, rng, want['total'] + 5)
    np.testing.assert_equal(want['codes'], codes)
    np.testing.assert_equal(want['sorted_ids'], order)
    np.testing.assert_equal(want['nodes'][1:], nodes[1:])
    np.testing.assert_equal(want['nodes'][['right_edge', 'data']][0], nodes[['right_edge', 'data']][0])
    np.testing.assert_equal(want['bounds'][:, :, :3], b[:, :, :3])
    assert want['total'] == total
    np.testing.assert_equal(want['pairs'], pairs)        # same ORDER too (sequential work-items)


@pytest.mark.parametrize("shape", G.AWKWARD_SHAPES)
def test_oracle_equals_reference_kernels_on_awkward_scenes(oracle, ref_env, shape):
    """golden_data.awkward_scene: duplicate codes, collinear / coplanar centres, radii from zero to scene-sized."""
    lib, dt = ref_env
    for seed in range(6):
        c, r = G.awkward_scene(shape, seed, dt)
        compare_with_reference_kernels(oracle, lib, dt, c, r)


def test_capacity_clamp(oracle, ref_env):
    lib, dt = ref_env
    c, r = G.random_spheres(500, dtype=dt)
    want = oracle.collide(c, r)
    *_, total, pairs = ref_pipeline(lib, dt, c, r, oracle.bounds(c), 7)
    assert total == want['total'] and np.array_equal(pairs, want['pairs'][:7])
    t2, p2, _ = oracle.traverse(want['nodes'], want['bounds'], capacity=7)
    assert t2 == total and np.array_equal(p2, pairs)


def test_golden_through_reference_kernels(oracle, ref_env):
    lib, dt = ref_env
    c = G.slots(G.SIX_COORDS, dt)
    rng = np.zeros((2, 4), dt)
    rng[0, :3] = G.SIX_COORDS.min(axis=0)
    rng[1, :3] = G.SIX_COORDS.max(axis=0)
    codes, order, nodes, b, total, pairs = ref_pipeline(lib, dt, c, np.ones(6, dt), rng, 2)
    np.testing.assert_equal(codes, G.SIX_CODES)
    assert total == 2 and G.canon(pairs) == G.SIX_PAIRS


def test_offsets_and_index(oracle):
    lib = oracle.ref("float32")
    if lib is None:
        pytest.skip("oracle/_ref not built")
    rs = np.random.RandomState(4)
    for vdt, fn in (("uint32", lib.ref_find_offsets_u32_u32), ("uint64", lib.ref_find_offsets_u64_u64)):
        v = np.sort(rs.randint(0, 300, 2000)).astype(vdt)
        n_off = 400
        out = np.full(n_off, len(v), vdt)          # host fill of collision/offset.py:43-46
        fn(C.c_size_t(len(v) - 1), _p(v), _p(out))
        np.testing.assert_equal(oracle.find_offsets(v, n_off, vdt), out)
    vals = rs.randint(0, 2**31, 500).astype(np.uint32)
    idx = rs.permutation(500).astype(np.uint32)
    g = np.zeros(500, np.uint32)
    lib.ref_gather_u32_u32(C.c_size_t(500), _p(vals), _p(idx), _p(g))
    np.testing.assert_equal(oracle.gather(vals, idx, np.zeros(500, np.uint32)), g)
    s = np.zeros(500, np.uint32)
    lib.ref_scatter_u32_u32(C.c_size_t(500), _p(vals), _p(idx), _p(s))
    np.testing.assert_equal(oracle.scatter(vals, idx, np.zeros(500, np.uint32)), s)
