"""CPU-only tests: host-side logic mirrored from the reference (sizing helpers,
validation errors, derived properties) and the C-ABI surface (the library loads
and exports every symbol include/collision_b200.h declares).  No compute calls."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
CTX = "cuda:0"     # a device handle only; nothing here touches the device


# ---------------------------------------------------------------- C ABI
def header_symbols():
    text = (ROOT / "include" / "collision_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"^\s*(?:int|size_t)\s+(col_\w+)\s*\(", text, flags=re.M)))


def test_library_exports_every_header_symbol():
    from collision_b200 import _lib, build
    build.build()
    handle = ctypes.CDLL(str(_lib.lib_path()))
    names = header_symbols()
    assert len(names) >= 20
    for name in names:
        assert hasattr(handle, name), name
    # the ctypes table binds exactly the header's surface
    assert sorted(_lib.SIGNATURES) == names
    L = _lib.lib()
    assert L.col_abi_version() == 2 and L.col_built_arch() == 100


def test_scratch_queries_are_pure_host():
    from collision_b200 import _lib
    L = _lib.lib()
    assert L.col_scan_scratch_bytes(1 << 20) >= (1 << 20) // 8192 * 8   # one 64-bit status word per 8192-value tile
    assert L.col_sort_scratch_bytes(1 << 20, 4, 4) > 2 * 4 * (1 << 20)
    assert L.col_sort_scratch_bytes(1 << 20, 2, 4) == 0
    assert L.col_reduce_scratch_bytes() > 0


def test_c_example_compiles_against_the_header():
    """examples/pipeline.c (the call sequence INTEGRATION.md shows a C host) type-checks against
    the header as C99 with warnings as errors, and links against the built library."""
    import subprocess
    import tempfile
    from collision_b200 import _lib
    src = ROOT / "examples" / "pipeline.c"
    with tempfile.TemporaryDirectory() as tmp:
        obj = Path(tmp) / "pipeline.so"
        subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-shared", "-fPIC", "-I", str(ROOT / "include"),
                        str(src), "-o", str(obj), "-L", str(_lib.lib_path().parent), "-l:" + _lib.lib_path().name,
                        "-Wl,--no-undefined", "-Wl,--allow-shlib-undefined"], check=True, capture_output=True)
        assert obj.exists()


def test_sm100a_only():
    import subprocess
    from collision_b200 import _lib
    out = subprocess.run(["cuobjdump", "-lelf", str(_lib.lib_path())], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_product_path_never_imports_oracle():
    for py in (ROOT / "collision_b200").glob("*.py"):
        assert "oracle" not in py.read_text(), py


# ---------------------------------------------------------------- misc
@pytest.mark.parametrize("x,base,expected", [(4, 5, 5), (5, 5, 5), (0, 5, 0), (4, 2, 4), (5, 2, 6), (0, 2, 0)])
def test_roundUp(x, base, expected):
    from collision_b200.misc import roundUp
    assert roundUp(x, base) == expected          # tests/test_misc.py:5-10


@pytest.mark.parametrize("x,expected", [(1, 1), (2, 2), (3, 4), (5, 8), (6, 8)])
def test_nextPowerOf2(x, expected):
    from collision_b200.misc import nextPowerOf2
    assert nextPowerOf2(x) == expected           # tests/test_misc.py:13-17


def test_product_and_dtypes():
    from collision_b200.misc import product, dtype_decl, dtype_sizeof
    assert product([1, 2, 3]) == 6 and product([]) == 1
    assert dtype_decl(np.dtype('uint32')) == 'uint'
    assert dtype_decl(np.dtype('float16')) == 'half'
    assert dtype_decl(np.dtype(('float16', 4))) == 'half4'
    with pytest.raises(ValueError):
        dtype_decl(np.dtype(('float16', 5)))
    with pytest.raises(ValueError):
        dtype_decl(np.dtype(('float16', (3, 4))))
    assert dtype_sizeof(np.dtype('uint32')) == 4
    assert dtype_sizeof(np.dtype('int64')) == 8
    assert dtype_sizeof(np.dtype(('float32', 3))) == 16
    assert dtype_sizeof(np.dtype(('float32', (4, 3)))) == 64
    assert dtype_sizeof(np.dtype((('float32', 3), 4))) == 64
    with pytest.raises(ValueError):
        dtype_sizeof(np.dtype(('float16', 5)))
    with pytest.raises(TypeError):
        dtype_sizeof(np.dtype([('foo', 'float32')]))
    with pytest.raises(TypeError):
        dtype_sizeof(np.dtype(([('foo', 'float32')], 4)))


# ---------------------------------------------------------------- scan
@pytest.mark.parametrize("size,group_size", [(1023, 4), (20, 4), (96, 6)])
def test_scanner_errs(size, group_size):
    from collision_b200.radix import PrefixScanner          # re-exported like upstream
    with pytest.raises(ValueError):
        PrefixScanner(CTX, size, group_size)                 # tests/test_scan_py.py:15-19


def test_scanner_resize_errs():
    from collision_b200.scan import PrefixScanner
    scanner = PrefixScanner(CTX, 1024, 4)
    with pytest.raises(ValueError):
        scanner.resize(1023, 4)
    assert (scanner.size, scanner.group_size) == (1024, 4)


@pytest.mark.parametrize("size,group_size,expected", [
    (1024, 4, (128, 16, 2)), (20, 2, (8, 2)), (24, 4, (8,)), (1032, 4, (136, 24, 4)), (160, 4, (24, 4)),
    (320, 4, (40, 8))])
def test_block_levels(size, group_size, expected):
    from collision_b200.scan import PrefixScanner
    assert PrefixScanner(CTX, size, group_size).block_lengths == expected   # tests/test_scan_py.py:32-45


# ---------------------------------------------------------------- radix
@pytest.mark.parametrize("size,group_size,bits", [(1024, 3, 4), (20, 4, 4), (1024, 4, 5), (1024, 4, 4), (64, 2, 4)])
def test_sorter_errs(size, group_size, bits):
    from collision_b200.radix import RadixSorter
    with pytest.raises(ValueError):                          # tests/test_radix_py.py:33-65 rules
        RadixSorter(CTX, size, group_size, bits)


def test_sorter_program_errs():
    from collision_b200.radix import RadixProgram, RadixSorter
    with pytest.raises(ValueError):
        RadixProgram(CTX, key_dtype=np.dtype('int32'))
    prog = RadixProgram(CTX, np.dtype('uint32'), np.dtype('float64'))
    with pytest.raises(ValueError):
        RadixSorter(CTX, 1024, 8, key_dtype=np.dtype('uint64'), value_dtype=np.dtype('float64'), program=prog)
    with pytest.raises(ValueError):
        RadixSorter(CTX, 1024, 8, key_dtype=np.dtype('uint32'), value_dtype=np.dtype('uint32'), program=prog)
    with pytest.raises(ValueError):
        RadixSorter("cuda:1", 1024, 8, value_dtype=np.dtype('float64'), program=prog)
    assert RadixSorter(CTX, 1024, 8, value_dtype=np.dtype('float64'), program=prog).program is prog


@pytest.mark.parametrize("kdt,bits,expected", [('uint32', 4, 8), ('uint32', 8, 4), ('uint64', 4, 16), ('uint32', 2, 16)])
def test_num_passes(kdt, bits, expected):
    from collision_b200.radix import RadixSorter
    s = RadixSorter(CTX, 1024, 128, bits, key_dtype=np.dtype(kdt))
    assert s.num_passes == expected                          # tests/test_radix_py.py:68-80


def test_histogram_len_and_resize():
    from collision_b200.radix import RadixSorter
    s = RadixSorter(CTX, 307200, 128)
    assert s.histogram_len == 19200
    s.resize(1 << 20, 128)
    assert s.histogram_len == 65536
    with pytest.raises(ValueError):
        s.resize(1000, 128)
    assert s.size == 1 << 20


# ---------------------------------------------------------------- collider
@pytest.mark.parametrize("size,ngroups,group_size,expected", [(48, 3, 8, 48), (47, 3, 8, 48), (49, 3, 8, 64)])
def test_padded_size(size, ngroups, group_size, expected):
    from collision_b200.collision import Collider
    c = Collider(CTX, size, ngroups, group_size, 'float32')
    assert c.padded_size == expected and c.n_nodes == 2 * size - 1   # tests/test_collision_py.py:40-46


@pytest.mark.parametrize("dt", ['float32', np.dtype('float32'), 'float64', np.dtype('float64')])
def test_collider_dtype(dt):
    from collision_b200.collision import Collider
    c = Collider(CTX, 100, 5, 8, coord_dtype=dt)
    assert c.program.coord_dtype == np.dtype(dt)              # tests/test_collision_py.py:329-336
    assert c.reducer.program.value_dtype == np.dtype((dt, 3))


def test_collider_errs_and_resize():
    from collision_b200.collision import Collider, CollisionProgram, Node, NO_NODE
    assert Node.itemsize == 16 and NO_NODE == 0xFFFFFFFF
    with pytest.raises(ValueError):
        CollisionProgram(CTX, 'int32')
    c = Collider(CTX, 350, 8, 64)
    with pytest.raises(ValueError):
        c.get_collisions(None, None, None, None, None, 5)     # collision.py:134-135
    c.resize(351, 8, 64)
    assert (c.size, c.padded_size, c.sorter.size) == (351, 384, 384)
    c.resize(size=1000)
    assert c.padded_size == 1024 and c.n_nodes == 1999
    assert (c.code_dtype, c.flag_dtype, c.counter_dtype, c.id_dtype) == (np.dtype('uint32'),) * 4


def test_index_offset_program_errs():
    from collision_b200.index import IndexProgram, Indexer
    from collision_b200.offset import OffsetProgram, OffsetFinder
    with pytest.raises(ValueError):
        IndexProgram(CTX, np.dtype('float32'), np.dtype('int32'))
    with pytest.raises(ValueError):
        OffsetProgram(CTX, np.dtype('int32'), np.dtype('uint32'))
    with pytest.raises(ValueError):
        OffsetProgram(CTX, np.dtype('uint32'), np.dtype('float32'))
    p = IndexProgram(CTX, np.dtype('float64'), np.dtype('uint64'))
    with pytest.raises(ValueError):
        Indexer(CTX, np.dtype('float64'), np.dtype('uint32'), program=p)
    q = OffsetProgram(CTX)
    with pytest.raises(ValueError):
        OffsetFinder(CTX, np.dtype('uint64'), np.dtype('uint32'), program=q)
    assert OffsetFinder(CTX, program=q).program is q


def test_bounds_program():
    from collision_b200.bounds import Bounds, BoundsProgram
    from collision_b200.summer import Summer, SumProgram
    p = BoundsProgram(CTX)
    assert p.value_dtype == np.dtype(('float32', 3)) and p.acc_dtype == np.dtype((('float32', 3), 2))
    with pytest.raises(ValueError):
        Bounds(CTX, 4, 8, np.dtype(('float64', 3)), program=p)
    r = Bounds(CTX, 4, 8, program=p)
    r.resize(8, 16)
    assert (r.ngroups, r.group_size) == (8, 16)
    assert SumProgram(CTX, np.dtype('float64')).acc_dtype == np.dtype(('float64', 1))
    assert Summer(CTX, 4, 8, np.dtype('float32')).program.value_dtype == np.dtype('float32')
    # integer sums (tests/test_summer_py.py:12); min/max stays floating point (its initialisers are +-INFINITY)
    assert SumProgram(CTX, np.dtype(('uint32', 4))).acc_dtype == np.dtype((('uint32', 4), 1))
    assert Summer(CTX, 4, 8, np.dtype('int32')).program.value_dtype == np.dtype('int32')
    with pytest.raises(ValueError):
        BoundsProgram(CTX, np.dtype(('uint32', 4)))
    with pytest.raises(ValueError):
        SumProgram(CTX, np.dtype('uint16'))
