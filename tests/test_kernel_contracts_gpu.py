"""The reference's KERNEL-LEVEL contracts of the non-collision programs, through
`program.kernels[name](cq, global_size, local_size, *args, g_times_l=...)` exactly as
upstream's tests launch them: tests/test_scan.py:24-104, tests/test_offset.py:31-47,
tests/test_radix.py:61-351 (block_sort, scatter, and the two chained into whole sorts),
collision/reduce.py:62-76 (bounds1 + bounds2 incl. the dropped-partial quirk), plus the
generic accumulator list of ReductionProgram (collision/reduce.py:9-22).
cl.LocalMemory arguments are placeholders here (None)."""
import numpy as np
import pytest

from gpu_util import dev_empty, to_dev, to_host

pytestmark = pytest.mark.gpu
LOCAL = None     # stands for cl.LocalMemory(...)


def prefix_sum(x):
    r = np.zeros_like(x)
    r[1:] = np.cumsum(x)[:-1]
    return r


def radix_key(values, radix_bits, radix_pass):
    return (values >> (radix_pass * radix_bits)) & ((2 ** radix_bits) - 1)


# ------------------------------------------------------------------- scan.cl
def test_scan_kernel(cuda_env):
    # tests/test_scan.py:24-60
    from collision_b200.scan import PrefixScanProgram
    dev, cq = cuda_env
    k = PrefixScanProgram(dev).kernels
    values = np.array([17, 6, 24, 28, 18, 22, 2, 1, 25, 17, 7, 17, 3, 19, 8, 23], dtype='uint32')
    block_size = 4
    nblocks = len(values) // 2 // block_size
    values_buf, sums_buf = to_dev(values), dev_empty(nblocks * 4)
    k['local_scan'](cq, (len(values) // 2,), (block_size,), values_buf, LOCAL, sums_buf).synchronize()
    np.testing.assert_equal(to_host(values_buf, 'uint32', values.shape),
                            [0, 17, 23, 47, 75, 93, 115, 117, 0, 25, 42, 49, 66, 69, 88, 96])
    np.testing.assert_equal(to_host(sums_buf, 'uint32', (nblocks,)), [118, 119])


def test_block_scan_kernel(cuda_env):
    # tests/test_scan.py:63-104
    from collision_b200.scan import PrefixScanProgram
    dev, cq = cuda_env
    k = PrefixScanProgram(dev).kernels
    values = np.array([0, 17, 23, 47, 75, 93, 115, 117, 0, 25, 42, 49, 66, 69, 88, 96], dtype='uint32')
    block_sums = np.array([118, 119], dtype='uint32')
    values_buf, sums_buf = to_dev(values), to_dev(block_sums)
    e = k['local_scan'](cq, (1,), (len(block_sums),), sums_buf, LOCAL, None, g_times_l=True)
    e.synchronize()
    np.testing.assert_equal(to_host(sums_buf, 'uint32', (2,)), [0, 118])
    k['block_scan'](cq, (len(values) // 2,), (4,), values_buf, sums_buf, wait_for=[e]).synchronize()
    np.testing.assert_equal(to_host(values_buf, 'uint32', values.shape),
                            [0, 17, 23, 47, 75, 93, 115, 117, 118, 143, 160, 167, 184, 187, 206, 214])


@pytest.mark.parametrize("size,group_size", [(1024, 4), (160, 4), (4096, 64), (1 << 16, 128)])
def test_scan_from_kernels(cuda_env, size, group_size):
    """PrefixScanner.prefix_sum rebuilt from the two kernels the way collision/scan.py:75-112 chains them."""
    from collision_b200.scan import PrefixScanner
    dev, cq = cuda_env
    sc = PrefixScanner(dev, size, group_size)
    k = sc.program.kernels
    rs = np.random.RandomState(4)
    values = rs.randint(0, 128, size).astype('uint32')
    buf = to_dev(values)
    lengths = sc.block_lengths
    sums = [dev_empty(l * 4, fill=0) for l in lengths]
    e = k['local_scan'](cq, (size // 2,), (group_size,), buf, LOCAL, sums[0])
    for l, a, b in zip(lengths, sums, sums[1:]):
        e = k['local_scan'](cq, (l // 2,), (group_size,), a, LOCAL, b, wait_for=[e])
    e = k['local_scan'](cq, (1,), (lengths[-1] // 2,), sums[-1], LOCAL, None, g_times_l=True, wait_for=[e])
    rev = list(reversed(sums))
    for l, a, b in zip(reversed(lengths[:-1]), rev, rev[1:]):
        e = k['block_scan'](cq, (l // 2,), (group_size,), b, a, wait_for=[e])
    k['block_scan'](cq, (size // 2,), (group_size,), buf, sums[0], wait_for=[e]).synchronize()
    np.testing.assert_equal(to_host(buf, 'uint32', (size,)), prefix_sum(values))


# ----------------------------------------------------------------- offset.cl
def test_offset_kernel(cuda_env):
    # tests/test_offset.py:31-47
    from collision_b200.offset import OffsetProgram
    dev, cq = cuda_env
    k = OffsetProgram(dev, np.dtype('uint32'), np.dtype('uint32')).kernels
    values = np.array([0, 0, 1, 1, 1, 1, 1, 3, 3, 3, 4, 5, 5], dtype='uint32')
    expected = np.array([0, 2, 7, 7, 10, 11], dtype='uint32')
    out = dev_empty(expected.nbytes, fill=0xEE)
    k['find_offsets'](cq, (len(values) - 1,), None, to_dev(values), out).synchronize()
    np.testing.assert_equal(to_host(out, 'uint32', expected.shape), expected)


# ------------------------------------------------------------------ radix.cl
SIZES = [(1, 8), (3, 8), (4, 8), (8, 32), (16, 128)]


@pytest.mark.parametrize("ngroups,group_size", SIZES)
@pytest.mark.parametrize("kdt", ["uint32", "uint64"])
def test_block_sort_random(cuda_env, kdt, ngroups, group_size):
    # tests/test_radix.py:61-118: every pass -- digit-major (H, ngroups) histogram == bincount,
    # every group == its stable argsort by the digit
    from collision_b200.radix import RadixProgram
    dev, cq = cuda_env
    kdt = np.dtype(kdt)
    k = RadixProgram(dev, kdt, np.dtype('uint32')).kernels
    radix_bits, H = 4, 16
    rs = np.random.RandomState(4)
    keys = rs.randint(0, 64, size=(ngroups, group_size * 2)).astype(kdt)
    hist_buf = dev_empty(ngroups * H * 4)
    for radix_pass in range(kdt.itemsize * 8 // radix_bits):
        keys_buf = to_dev(keys)
        k['block_sort'](cq, (ngroups,), (group_size,), keys_buf, LOCAL, LOCAL, None, LOCAL, LOCAL,
                        hist_buf, LOCAL, LOCAL, radix_bits, radix_pass, g_times_l=True).synchronize()
        digits = radix_key(keys, radix_bits, radix_pass)
        order = np.argsort(digits, kind='mergesort', axis=1)
        hist = to_host(hist_buf, 'uint32', (H, ngroups))
        for g in range(ngroups):
            np.testing.assert_equal(hist[:, g], np.bincount(digits[g].astype('int64'), minlength=H))
        np.testing.assert_equal(to_host(keys_buf, kdt, keys.shape), np.take_along_axis(keys, order, axis=1))


@pytest.mark.parametrize("ngroups,group_size", SIZES)
@pytest.mark.parametrize("kdt", ["uint32", "uint64"])
def test_scatter_kernel(cuda_env, kdt, ngroups, group_size):
    # tests/test_radix.py:122-182
    from collision_b200.radix import RadixProgram
    dev, cq = cuda_env
    kdt = np.dtype(kdt)
    k = RadixProgram(dev, kdt, np.dtype('uint32')).kernels
    radix_bits, H = 4, 16
    rs = np.random.RandomState(4)
    keys = rs.randint(0, 64, size=(ngroups, group_size * 2)).astype(kdt)
    for radix_pass in range(kdt.itemsize * 8 // radix_bits):
        digits = radix_key(keys, radix_bits, radix_pass)
        block_keys = np.take_along_axis(keys, np.argsort(digits, kind='mergesort', axis=1), axis=1)
        digits = radix_key(block_keys, radix_bits, radix_pass).astype('int64')
        hist = np.array([np.bincount(g, minlength=H) for g in digits], dtype='uint32').T.copy()
        offset = prefix_sum(hist.reshape(-1)).reshape(H, ngroups)
        out = dev_empty(keys.nbytes)
        k['scatter'](cq, (ngroups,), (group_size,), to_dev(block_keys), out, None, None,
                     to_dev(offset), LOCAL, to_dev(hist), LOCAL, radix_bits, radix_pass, g_times_l=True).synchronize()
        expected = block_keys.reshape(-1)[np.argsort(digits, axis=None, kind='mergesort')]
        np.testing.assert_equal(to_host(out, kdt, keys.shape), expected.reshape(keys.shape))


@pytest.mark.parametrize("ngroups,group_size", SIZES)
@pytest.mark.parametrize("kdt", ["uint32", "uint64"])
@pytest.mark.parametrize("vdt", [None, "uint32", "float64"])
def test_sort_from_kernels(cuda_env, kdt, vdt, ngroups, group_size):
    # tests/test_radix.py:185-351: block_sort -> scan of the digit-major histogram -> scatter ->
    # copy back, pass after pass; after every pass == numpy's stable pass, at the end == stable argsort
    from collision_b200.radix import RadixProgram
    dev, cq = cuda_env
    kdt = np.dtype(kdt)
    vdt = np.dtype(vdt) if vdt else None
    k = RadixProgram(dev, kdt, vdt or np.dtype('uint32')).kernels
    radix_bits, H = 4, 16
    rs = np.random.RandomState(4)
    size = ngroups * group_size * 2
    keys = rs.randint(0, 1 << 16, size).astype(kdt)
    values = None if vdt is None else rs.random_sample(size).astype(vdt) if vdt.kind == 'f' else np.arange(size, dtype=vdt)
    cur_k, cur_v = keys.copy(), None if values is None else values.copy()
    for radix_pass in range(kdt.itemsize * 8 // radix_bits):
        keys_buf, out_keys = to_dev(cur_k), dev_empty(cur_k.nbytes)
        vals_buf = None if cur_v is None else to_dev(cur_v)
        out_vals = None if cur_v is None else dev_empty(cur_v.nbytes)
        hist_buf = dev_empty(ngroups * H * 4)
        e = k['block_sort'](cq, (size // 2,), (group_size,), keys_buf, LOCAL, LOCAL, vals_buf, LOCAL, LOCAL,
                            hist_buf, LOCAL, LOCAL, radix_bits, radix_pass)
        e.synchronize()
        offset = prefix_sum(to_host(hist_buf, 'uint32', (H * ngroups,)))
        k['scatter'](cq, (size // 2,), (group_size,), keys_buf, out_keys, vals_buf, out_vals,
                     to_dev(offset), LOCAL, hist_buf, LOCAL, radix_bits, radix_pass).synchronize()
        order = np.argsort(radix_key(cur_k, radix_bits, radix_pass), kind='mergesort')
        cur_k = cur_k[order]
        np.testing.assert_equal(to_host(out_keys, kdt, (size,)), cur_k)
        if cur_v is not None:
            cur_v = cur_v[order]
            np.testing.assert_equal(to_host(out_vals, vdt, (size,)), cur_v)
    order = np.argsort(keys, kind='mergesort')
    np.testing.assert_equal(cur_k, keys[order])
    if values is not None:
        np.testing.assert_equal(cur_v, values[order])


# ----------------------------------------------------------------- reduce.cl
@pytest.mark.parametrize("size,ngroups,group_size", [(24, 2, 4), (100, 4, 8), (3000, 3, 8), (3000, 5, 16), (3000, 10, 32)])
@pytest.mark.parametrize("coord_dtype", [("float32", 3), ("float64", 4), "float32"])
def test_bounds_kernels_reproduce_the_quirk(cuda_env, oracle, coord_dtype, size, ngroups, group_size):
    """bounds1 + bounds2 with the reference's geometry: the per-group partials, and the result
    INCLUDING upstream's dropped tail partials for ngroups 3, 5, 10 (reduce.cl:42-52),
    against the oracle's emulation of that launch."""
    from collision_b200.bounds import BoundsProgram
    from collision_b200.misc import dtype_lanes
    dev, cq = cuda_env
    dt = np.dtype(coord_dtype)
    lanes = dtype_lanes(dt)
    k = BoundsProgram(dev, dt).kernels
    rs = np.random.RandomState(4)
    vals = (rs.random_sample((size, lanes)) * 100 - 50).astype(dt.base)
    groups = dev_empty(ngroups * 2 * lanes * dt.base.itemsize)
    out = dev_empty(2 * lanes * dt.base.itemsize)
    e = k['bounds1'](cq, (ngroups,), (group_size,), to_dev(vals), size, groups, LOCAL, g_times_l=True)
    e.synchronize()
    part = to_host(groups, dt.base, (ngroups, 2, lanes)).copy()
    k['bounds2'](cq, (1,), (ngroups,), groups, out, g_times_l=True, wait_for=[e]).synchronize()
    got = to_host(out, dt.base, (2, lanes))
    use = slice(0, 3) if lanes == 4 and dt.shape == (3,) else slice(None)
    want = oracle.bounds(vals, ngroups, group_size)
    np.testing.assert_equal(got[:, use], want[:, use])
    # the partials: group g holds the min/max of the items its work-items stride over
    for g in range(ngroups):
        idx = np.concatenate([np.arange(g * group_size + t, size, ngroups * group_size) for t in range(group_size)])
        if len(idx):
            np.testing.assert_equal(part[g, 0, use], vals[idx].min(axis=0)[use])
            np.testing.assert_equal(part[g, 1, use], vals[idx].max(axis=0)[use])
    if ngroups & (ngroups - 1) == 0:
        np.testing.assert_equal(got[0, use], vals.min(axis=0)[use])
        np.testing.assert_equal(got[1, use], vals.max(axis=0)[use])


def test_generic_accumulator_list(cuda_env):
    """collision/reduce.py:9-22: a ReductionProgram subclass with any accumulator list builds
    and reduces (here: sum, min and max at once), through the same Reducer.reduce call."""
    from collision_b200.reduce import ReductionProgram, Reducer
    dev, cq = cuda_env

    class Stats(ReductionProgram):
        accumulator = [("0", "ADD"), ("INFINITY", "min"), ("-INFINITY", "max")]

    class StatsReducer(Reducer):
        program_type = Stats

    for dt, lanes in (("float64", 4), ("int32", 1), ("uint32", 4)):
        vdt = np.dtype((dt, lanes)) if lanes > 1 else np.dtype(dt)
        rs = np.random.RandomState(7)
        size = 5000
        vals = rs.randint(-1000, 1000, (size, lanes)).astype(dt) if dt != "uint32" else \
            rs.randint(0, 1 << 31, (size, lanes)).astype(dt)
        red = StatsReducer(dev, 8, 64, vdt)
        assert red.program.acc_dtype == np.dtype((vdt, 3))
        out = dev_empty(3 * lanes * np.dtype(dt).itemsize)
        red.reduce(cq, size, to_dev(vals), out).synchronize()
        got = to_host(out, dt, (3, lanes))
        np.testing.assert_equal(got[0], vals.sum(axis=0, dtype=dt))       # integer sums wrap; these f64 sums are exact
        np.testing.assert_equal(got[1], vals.min(axis=0))
        np.testing.assert_equal(got[2], vals.max(axis=0))

    class Bad(ReductionProgram):
        accumulator = [("0", "mul")]
    with pytest.raises(ValueError):
        Bad(dev, np.dtype('float32'))


# ------------------------------------------------------------------ index.cl
def test_index_kernels(cuda_env):
    from collision_b200.index import IndexProgram
    dev, cq = cuda_env
    k = IndexProgram(dev, np.dtype(('float32', 3)), np.dtype('uint32')).kernels
    rs = np.random.RandomState(4)
    n = 1000
    vals = rs.random_sample((n, 4)).astype('float32')
    idx = rs.permutation(n).astype('uint32')
    out = dev_empty(vals.nbytes)
    k['gather'](cq, (n,), None, to_dev(vals), to_dev(idx), out).synchronize()
    np.testing.assert_equal(to_host(out, 'float32', (n, 4)), vals[idx])
    out = dev_empty(vals.nbytes)
    k['scatter'](cq, (n,), None, to_dev(vals), to_dev(idx), out).synchronize()
    want = np.empty_like(vals)
    want[idx] = vals
    np.testing.assert_equal(to_host(out, 'float32', (n, 4)), want)


# -------------------------------------------------------------- collision.cl
def test_range_kernel(cuda_env):
    # collision/collision.cl:8-10 through kernels['range'] (collision/collision.py:143-146)
    from collision_b200.collision import CollisionProgram
    dev, cq = cuda_env
    k = CollisionProgram(dev).kernels
    for n in (1, 48, 4097):
        buf = dev_empty((n + 3) * 4, fill=0xAB)
        k['range'](cq, (n,), None, buf).synchronize()
        got = to_host(buf, 'uint32', (n + 3,))
        np.testing.assert_equal(got[:n], np.arange(n))
        assert (got[n:] == 0xABABABAB).all()
