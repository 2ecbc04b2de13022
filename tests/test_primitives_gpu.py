"""GPU parity tests of the standalone primitives (sort, scan, bounds, sum, offsets,
gather/scatter) against the CPU oracle / numpy, mirroring upstream's
tests/test_radix*.py, test_scan*.py, test_bounds_py.py, test_summer_py.py,
test_offset*.py and test_index.py."""
import numpy as np
import pytest

import golden_data as G
from gpu_util import dev_empty, to_dev, to_host

pytestmark = pytest.mark.gpu


# -------------------------------------------------------------------- scan
def test_scan_golden(cuda_env):
    # tests/test_scan.py:24-104
    from collision_b200.scan import PrefixScanner
    dev, cq = cuda_env
    values = np.array([17, 6, 24, 28, 18, 22, 2, 1, 25, 17, 7, 17, 3, 19, 8, 23], dtype='uint32')
    full = np.array([0, 17, 23, 47, 75, 93, 115, 117, 118, 143, 160, 167, 184, 187, 206, 214], dtype='uint32')
    buf = to_dev(values)
    PrefixScanner(dev, 16, 4).prefix_sum(cq, buf).synchronize()
    np.testing.assert_equal(to_host(buf, 'uint32', (16,)), full)


@pytest.mark.parametrize("size,group_size", [(20, 2), (24, 4), (1024, 4), (160, 4), (320, 4), (4096, 8),
                                             (4096 * 37 + 256, 128), (1 << 22, 128)])
def test_prefix_sum(cuda_env, oracle, size, group_size):
    # tests/test_scan_py.py:47-64; multi-tile sizes exercise the look-back chain
    from collision_b200.scan import PrefixScanner
    dev, cq = cuda_env
    scanner = PrefixScanner(dev, size, group_size)
    rs = np.random.RandomState(4)
    for hi in (size, 2**32):           # small values and values that wrap mod 2^32
        values = rs.randint(0, hi, size=size, dtype=np.uint64).astype('uint32')
        buf = to_dev(values)
        scanner.prefix_sum(cq, buf).synchronize()
        np.testing.assert_equal(to_host(buf, 'uint32', (size,)), oracle.exclusive_scan(values))


@pytest.mark.parametrize("size", [1 << 23, (1 << 23) + 3, (1 << 24) + 4097, (1 << 25) + 1, 3 * (1 << 23) + 777])
@pytest.mark.parametrize("misaligned", [False, True])
def test_scan_large(cuda_env, size, misaligned):
    """From 2^23 values on the scan runs as a persistent chunked grid (sum a tile, scan it two chunks
    later out of L2): ragged last chunk / last tile, wrapping values, the u64 grand total, and the
    fall-back to the single-pass kernel when the buffer is not 16-byte aligned."""
    import torch
    from collision_b200 import _lib
    dev, cq = cuda_env
    L, P, S = _lib.lib(), _lib.ptr, cq.cuda_stream
    g = torch.Generator(device=dev).manual_seed(size & 0xFFFF)
    wide = torch.randint(0, 1 << 32, (size,), dtype=torch.int64, device=dev, generator=g)
    store = torch.empty(size + 1, dtype=torch.int32, device=dev)
    data = store[1:] if misaligned else store[:size]
    data.copy_(wide.to(torch.int32))
    want = torch.cumsum(wide, 0)
    scratch = _lib.alloc_bytes(dev, L.col_scan_scratch_bytes(size))
    tot = torch.zeros(1, dtype=torch.int64, device=dev)
    cq.wait_stream(torch.cuda.current_stream(dev))
    for _ in range(2):      # the scratch is reusable as it is left
        data.copy_(wide.to(torch.int32))
        cq.wait_stream(torch.cuda.current_stream(dev))
        assert L.col_exclusive_scan_u32(data.data_ptr(), size, P(tot), P(scratch), S) == 0
        cq.synchronize()
        assert int(tot.item()) == int(want[-1].item())
        assert data[0].item() == 0
        assert torch.equal(data[1:].to(torch.int64) & 0xFFFFFFFF, want[:-1] & 0xFFFFFFFF)


def test_scan_large_in_a_cuda_graph(cuda_env):
    """The large-array scan is a cooperative launch; it must capture into a CUDA graph and replay
    (Collider(graph=True) at 2^24 spheres and more goes through it)."""
    import torch
    from collision_b200 import _lib
    dev, cq = cuda_env
    L, P = _lib.lib(), _lib.ptr
    size = (1 << 24) + 5000
    g = torch.Generator(device=dev).manual_seed(11)
    src = torch.randint(0, 1000, (size,), dtype=torch.int32, device=dev, generator=g)
    want = torch.cumsum(src.to(torch.int64), 0)
    data = torch.empty_like(src)
    scratch = _lib.alloc_bytes(dev, L.col_scan_scratch_bytes(size))
    tot = torch.zeros(1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize(dev)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=cq):
        data.copy_(src)
        assert L.col_exclusive_scan_u32(P(data), size, P(tot), P(scratch), torch.cuda.current_stream(dev).cuda_stream) == 0
    for _ in range(2):
        data.zero_()
        tot.zero_()
        graph.replay()
        torch.cuda.synchronize(dev)
        assert int(tot.item()) == int(want[-1].item())
        assert data[0].item() == 0 and torch.equal(data[1:].to(torch.int64) & 0xFFFFFFFF, want[:-1] & 0xFFFFFFFF)


def test_scanner_resize(cuda_env, oracle):
    # tests/test_scan_py.py:67-92
    from collision_b200.scan import PrefixScanner
    dev, cq = cuda_env
    scanner = PrefixScanner(dev, 1024, 4)
    scanner.resize(1 << 16, 8)
    values = np.random.RandomState(4).randint(0, 100, size=1 << 16).astype('uint32')
    buf = to_dev(values)
    scanner.prefix_sum(cq, buf).synchronize()
    np.testing.assert_equal(to_host(buf, 'uint32', (1 << 16,)), oracle.exclusive_scan(values))


# -------------------------------------------------------------------- sort
def make_keys(rs, n, kdt, kind):
    if kind == "small":
        return rs.randint(0, 1000, size=n).astype(kdt)              # tests/benchmarks/test_radix.py:52
    if kind == "sorted":
        return np.arange(n).astype(kdt)
    if kind == "dups":
        return (rs.randint(0, 4, size=n) * 0x01010101).astype(kdt)
    hi = 2**32 if np.dtype(kdt).itemsize == 4 else 2**63
    return rs.randint(0, hi, size=n, dtype=np.uint64).astype(kdt)


def make_values(rs, n, vdt):
    vdt = np.dtype(vdt)
    lanes = (4,) if vdt.shape in ((3,), (4,)) else ()
    return (rs.random_sample((n,) + lanes) * 1e6).astype(vdt.base)


@pytest.mark.parametrize("size,group_size", [(32, 8), (15360, 32), (32, 16), (8192 * 3 + 512, 128), (1 << 20, 128)])
@pytest.mark.parametrize("kdt", ["uint32", "uint64"])
@pytest.mark.parametrize("kind", ["random", "small", "sorted", "dups"])
def test_sorter_keys(cuda_env, oracle, size, group_size, kdt, kind):
    # tests/test_radix_py.py:83-136
    from collision_b200.radix import RadixSorter
    dev, cq = cuda_env
    keys = make_keys(np.random.RandomState(4), size, kdt, kind)
    sorter = RadixSorter(dev, size, group_size, key_dtype=np.dtype(kdt))
    kin, kout = to_dev(keys), dev_empty(keys.nbytes)
    sorter.sort(cq, kin, kout).synchronize()
    np.testing.assert_equal(to_host(kout, kdt, (size,)), np.sort(keys, kind='stable'))
    np.testing.assert_equal(to_host(kin, kdt, (size,)), keys)       # input left untouched


@pytest.mark.parametrize("size,group_size", [(32, 8), (15360, 32), (8192 * 5 + 1024, 128), (1 << 20, 128)])
@pytest.mark.parametrize("kdt", ["uint32", "uint64"])
@pytest.mark.parametrize("vdt", ["uint32", "float64", ("float64", 3), ("float64", 4), ("float32", 3)])
def test_arg_sorter(cuda_env, oracle, size, group_size, kdt, vdt):
    # tests/test_radix_py.py:139-201: values[argsort(keys, mergesort)] -- stability asserted
    from collision_b200.radix import RadixSorter
    dev, cq = cuda_env
    rs = np.random.RandomState(4)
    keys = make_keys(rs, size, kdt, "small")          # many duplicates
    values = make_values(rs, size, vdt)
    sorter = RadixSorter(dev, size, group_size, key_dtype=np.dtype(kdt), value_dtype=np.dtype(vdt))
    kout, vout = dev_empty(keys.nbytes), dev_empty(values.nbytes)
    sorter.sort(cq, to_dev(keys), kout, to_dev(values), vout).synchronize()
    ko, vo = oracle.sort_pairs(keys, values)
    np.testing.assert_equal(to_host(kout, kdt, (size,)), ko)
    np.testing.assert_equal(to_host(vout, values.dtype, values.shape), vo)
    order = np.argsort(keys, kind='mergesort')
    np.testing.assert_equal(vo, values[order])


@pytest.mark.parametrize("kdt,radix_bits", [("uint32", 4), ("uint32", 8), ("uint64", 4), ("uint32", 2)])
def test_every_pass(cuda_env, oracle, kdt, radix_bits):
    # tests/test_radix.py:122-253: after EVERY pass keys/values == the stable partition by that digit
    from collision_b200.radix import RadixSorter
    dev, cq = cuda_env
    size = 8192 * 2 + 256
    rs = np.random.RandomState(4)
    keys = make_keys(rs, size, kdt, "random")
    values = rs.randint(0, 2**31, size=size).astype('uint32')
    sorter = RadixSorter(dev, size, 128, radix_bits=radix_bits, key_dtype=np.dtype(kdt))
    assert sorter.num_passes == np.dtype(kdt).itemsize * 8 // radix_bits
    k, v = keys, values
    for p in range(sorter.num_passes):
        kout, vout = dev_empty(k.nbytes), dev_empty(v.nbytes)
        sorter.radix_pass(cq, to_dev(k), kout, to_dev(v), vout, p).synchronize()
        k2, v2 = oracle.radix_pass(k, v, radix_bits, p)
        np.testing.assert_equal(to_host(kout, kdt, (size,)), k2)
        np.testing.assert_equal(to_host(vout, 'uint32', (size,)), v2)
        k, v = k2, v2
    np.testing.assert_equal(k, np.sort(keys, kind='stable'))


def test_sorter_resize(cuda_env):
    # tests/test_radix_py.py:108-136
    from collision_b200.radix import RadixSorter
    dev, cq = cuda_env
    sorter = RadixSorter(dev, 32, 8)
    sorter.resize(15360, 32)
    keys = np.random.RandomState(4).randint(0, 2**31, size=15360).astype('uint32')
    kout = dev_empty(keys.nbytes)
    sorter.sort(cq, to_dev(keys), kout).synchronize()
    np.testing.assert_equal(to_host(kout, 'uint32', (15360,)), np.sort(keys))


# ------------------------------------------------------------------ bounds
@pytest.mark.parametrize("coord_dtype", [("float32", 3), ("float64", 4), "float32", "float64", ("float32", 2),
                                         ("float32", 8), ("float64", 16)])
@pytest.mark.parametrize("size,ngroups,group_size", [(24, 2, 4), (100, 4, 8), (3000, 3, 8), (1 << 20, 64, 128)])
def test_bounds(cuda_env, oracle, coord_dtype, size, ngroups, group_size):
    # tests/test_bounds_py.py:18-148 (normal, negative-only, 1e8-scaled data)
    from collision_b200.bounds import Bounds
    from collision_b200.misc import dtype_sizeof
    dev, cq = cuda_env
    cdt = np.dtype(coord_dtype)
    lanes = dtype_sizeof(cdt) // cdt.base.itemsize
    rs = np.random.RandomState(4)
    reducer = Bounds(dev, ngroups, group_size, cdt)
    for gen in (lambda: rs.normal(size=(size, lanes)),
                lambda: rs.normal(-10, 1, size=(size, lanes)).clip(max=-1.0),
                lambda: rs.normal(size=(size, lanes)) * 1e8):
        values = gen().astype(cdt.base)
        out_buf = dev_empty(2 * dtype_sizeof(cdt))
        reducer.reduce(cq, size, to_dev(values), out_buf).synchronize()
        got = to_host(out_buf, cdt.base, (2, lanes))
        np.testing.assert_equal(got, np.stack([values.min(axis=0), values.max(axis=0)]))
        np.testing.assert_equal(got, oracle.bounds(values))


def test_bounds_ignores_pad_lane(cuda_env):
    from collision_b200.bounds import Bounds
    dev, cq = cuda_env
    c, _ = G.random_spheres(10000, pad=np.nan)
    out_buf = dev_empty(32)
    Bounds(dev, 8, 128).reduce(cq, len(c), to_dev(c), out_buf).synchronize()
    got = to_host(out_buf, 'float32', (2, 4))
    np.testing.assert_equal(got[0, :3], c[:, :3].min(axis=0))
    np.testing.assert_equal(got[1, :3], c[:, :3].max(axis=0))


@pytest.mark.parametrize("coord_dtype", [("float64", 3), "float32", ("float64", 4)])
def test_sum(cuda_env, coord_dtype):
    # tests/test_summer_py.py:20-53 (floating dtypes; compared almost-equal as upstream does)
    from collision_b200.summer import Summer
    from collision_b200.misc import dtype_sizeof
    dev, cq = cuda_env
    cdt = np.dtype(coord_dtype)
    lanes = dtype_sizeof(cdt) // cdt.base.itemsize
    values = np.random.RandomState(4).uniform(0, 200, size=(100, lanes)).astype(cdt.base)
    out_buf = dev_empty(dtype_sizeof(cdt))
    Summer(dev, 4, 8, cdt).reduce(cq, 100, to_dev(values), out_buf).synchronize()
    got = to_host(out_buf, cdt.base, (lanes,))
    np.testing.assert_allclose(got[:3], values.sum(axis=0)[:3], rtol=1e-5)


@pytest.mark.parametrize("size", [100, 1000003])
@pytest.mark.parametrize("coord_dtype", [("uint32", 4), "int32", ("uint32", 3), "uint64", ("int64", 2)])
def test_sum_integer(cuda_env, coord_dtype, size):
    # tests/test_summer_py.py:12 (uint32x4, int32): integer sums wrap, so they are exact in any order
    from collision_b200.summer import Summer
    from collision_b200.misc import dtype_sizeof
    dev, cq = cuda_env
    cdt = np.dtype(coord_dtype)
    lanes = dtype_sizeof(cdt) // cdt.base.itemsize
    rng = np.random.RandomState(5)
    if size == 100:
        values = rng.uniform(0, 200, size=(size, lanes)).astype(cdt.base)       # upstream's values
    else:
        info = np.iinfo(cdt.base)                                                 # full range: the sum wraps
        values = rng.randint(info.min, info.max, size=(size, lanes), dtype=cdt.base)
    out_buf = dev_empty(dtype_sizeof(cdt), fill=0xEE)
    Summer(dev, 4, 8, cdt).reduce(cq, size, to_dev(values), out_buf).synchronize()
    got = to_host(out_buf, cdt.base, (lanes,))
    with np.errstate(over='ignore'):
        np.testing.assert_equal(got, values.sum(axis=0, dtype=cdt.base))


# ----------------------------------------------------------------- offsets
@pytest.mark.parametrize("vdt,odt", [("uint32", "uint32"), ("uint64", "uint64"), ("uint32", "uint64")])
def test_offsets(cuda_env, oracle, vdt, odt):
    # tests/test_offset.py:31-47, tests/test_offset_py.py:23-62
    from collision_b200.offset import OffsetFinder
    dev, cq = cuda_env
    finder = OffsetFinder(dev, np.dtype(vdt), np.dtype(odt))

    def run(values, n_offsets):
        out = dev_empty(n_offsets * np.dtype(odt).itemsize, fill=0xEE)
        finder.find_offsets(cq, to_dev(values.astype(vdt)), len(values), out, n_offsets).synchronize()
        return to_host(out, odt, (n_offsets,))

    np.testing.assert_equal(run(G.OFFSET_VALUES, 7), G.OFFSET_EXPECTED)
    missing = np.array([1, 1, 1, 1, 1, 1, 1, 3, 3, 3, 3, 3, 3])
    np.testing.assert_equal(run(missing, 7), [0, 0, 7, 7, 13, 13, 13])
    rs = np.random.RandomState(4)
    for maxval in (2000, 2000000):          # tests/benchmarks/test_offset.py:24-39 shapes (scaled)
        v = np.sort(rs.randint(0, maxval, size=1 << 16))
        np.testing.assert_equal(run(v, maxval + 10), oracle.find_offsets(v.astype(vdt), maxval + 10, odt))
    # the reference's corner: entries v <= values[0] with v >= n_values-1 keep the fill
    v = np.array([5, 5, 6])
    np.testing.assert_equal(run(v, 8), oracle.find_offsets(v.astype(vdt), 8, odt))


# ---------------------------------------------------------- gather/scatter
@pytest.mark.parametrize("vdt", ["uint32", "float64", ("float32", 3), ("float64", 4), "uint8"])
@pytest.mark.parametrize("idt", ["uint32", "uint64"])
def test_gather_scatter(cuda_env, oracle, vdt, idt):
    # tests/test_index.py:18-79
    from collision_b200.index import Indexer
    from collision_b200.misc import dtype_sizeof
    dev, cq = cuda_env
    vdt = np.dtype(vdt)
    lanes = dtype_sizeof(vdt) // vdt.base.itemsize
    rs = np.random.RandomState(4)
    n_in, n = 5000, 3000
    shape = (lanes,) if lanes > 1 else ()
    vals = (rs.random_sample((n_in,) + shape) * 200).astype(vdt.base)
    idx = rs.permutation(n_in)[:n].astype(idt)
    indexer = Indexer(dev, vdt, np.dtype(idt))
    out = dev_empty(n * dtype_sizeof(vdt), fill=0)
    indexer.gather(cq, n, to_dev(vals), to_dev(idx), out).synchronize()
    np.testing.assert_equal(to_host(out, vdt.base, (n,) + shape), vals[idx])
    fill = np.zeros((n_in,) + shape, vdt.base)
    out = to_dev(fill)
    indexer.scatter(cq, n, to_dev(vals[:n]), to_dev(idx), out).synchronize()
    exp = fill.copy()
    exp[idx] = vals[:n]
    np.testing.assert_equal(to_host(out, vdt.base, (n_in,) + shape), exp)   # untouched slots keep the fill
