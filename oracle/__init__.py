"""CPU ORACLE -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

numpy/ctypes face of ``oracle/collision_oracle.c`` (a plain-C restatement of
the reference's OpenCL kernels; every C function cites the reference
file:line it follows) and of ``oracle/_ref/libref_*.so`` (the reference's own
barrier-free ``.cl`` kernels compiled as C++ where the reference checkout is
available).  Parity status: PINNED (see the header of collision_oracle.c).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` leg may import this package.  ``collision_b200`` never does.
"""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = None

Node = np.dtype([('parent', 'uint32'), ('right_edge', 'uint32'), ('data', 'uint32', 2)])
NO_NODE = 0xFFFFFFFF


def _cpu_stamp():
    """Identifies the host ISA the library was compiled for (-march=native)."""
    import hashlib
    flags = ""
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith(("flags", "Features", "model name")):
                flags += line
                if line.startswith(("flags", "Features")):
                    break
    except OSError:
        pass
    return hashlib.sha1(flags.encode()).hexdigest()[:16]


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference exists) with
    -O3 -march=native; rebuilt when the source is newer or when the host CPU is not the
    one the library was built on (the built file travels to the GPU box)."""
    so = _HERE / "liboracle.so"
    src = _HERE / "collision_oracle.c"
    stamp = _HERE / "liboracle.stamp"
    want = _cpu_stamp()
    have = stamp.read_text().strip() if stamp.exists() else ""
    stale = not so.exists() or so.stat().st_mtime < max(src.stat().st_mtime, (_HERE / "Makefile").stat().st_mtime)
    if force or stale or have != want:
        subprocess.run(["make", "-C", str(_HERE), "-s", "-B", "all"], check=True)
        stamp.write_text(want + "\n")
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = _HERE / "liboracle.so"
        if not so.exists():
            build()
        _LIB = C.CDLL(str(so))
        _LIB.orc_traverse_f32.restype = C.c_uint64
        _LIB.orc_traverse_f64.restype = C.c_uint64
        _LIB.orc_traverse_f32_mt.restype = C.c_uint64
        _LIB.orc_brute_force_f32.restype = C.c_uint64
        _LIB.orc_brute_force_f64.restype = C.c_uint64
        _LIB.orc_pipeline_f32.restype = C.c_uint64
        _LIB.orc_pipeline_f32_single.restype = C.c_uint64
        _LIB.orc_traverse_f32_single.restype = C.c_uint64
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _sfx(dt):
    dt = np.dtype(dt)
    if dt == np.float32:
        return "f32"
    if dt == np.float64:
        return "f64"
    raise ValueError(dt)


def num_threads():
    return lib().orc_num_threads()


def set_num_threads(nt=None):
    """Use `nt` OpenMP threads (default: every core this process may run on), whatever
    OMP_NUM_THREADS says -- torchrun exports OMP_NUM_THREADS=1.  Returns the count in use."""
    import os
    if nt is None:
        nt = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    lib().orc_set_num_threads(C.c_int(int(nt)))
    return num_threads()


# ------------------------------------------------------------------ stages
def bounds(values, ngroups=0, group_size=0):
    """values: (n, lanes) float32/float64, C-contiguous.  Returns (2, lanes).
    ngroups/group_size != 0 emulates the reference launch geometry including
    the dropped-partial quirk (quirk_ngroups switch of SURVEY 0.1-4)."""
    v = np.ascontiguousarray(values)
    if v.ndim == 1:
        v = v[:, None]
    n, lanes = v.shape
    out = np.empty((2, lanes), v.dtype)
    getattr(lib(), "orc_bounds_" + _sfx(v.dtype))(
        _p(v), C.c_uint64(n), C.c_uint32(lanes), _p(out), C.c_uint32(ngroups), C.c_uint32(group_size))
    return out


def morton(coords4, range2x4):
    c = np.ascontiguousarray(coords4)
    r = np.ascontiguousarray(range2x4, dtype=c.dtype)
    assert c.ndim == 2 and c.shape[1] == 4 and r.shape == (2, 4)
    codes = np.empty(len(c), np.uint32)
    getattr(lib(), "orc_morton_" + _sfx(c.dtype))(_p(codes), _p(c), _p(r), C.c_uint32(len(c)))
    return codes


def _value_bytes(values):
    return 0 if values is None else values.dtype.itemsize * int(np.prod(values.shape[1:], dtype=np.int64))


def radix_pass(keys, values, radix_bits, pass_):
    k = np.ascontiguousarray(keys)
    v = None if values is None else np.ascontiguousarray(values)
    ko = np.empty_like(k)
    vo = None if v is None else np.empty_like(v)
    lib().orc_radix_pass(_p(k), _p(ko), _p(v), _p(vo), C.c_uint64(len(k)), C.c_uint32(k.dtype.itemsize),
                         C.c_uint32(_value_bytes(v)), C.c_uint32(radix_bits), C.c_uint32(pass_))
    return ko, vo


def sort_pairs(keys, values=None):
    k = np.ascontiguousarray(keys)
    v = None if values is None else np.ascontiguousarray(values)
    ko = np.empty_like(k)
    vo = None if v is None else np.empty_like(v)
    lib().orc_sort_pairs(_p(k), _p(ko), _p(v), _p(vo), C.c_uint64(len(k)), C.c_uint32(k.dtype.itemsize),
                         C.c_uint32(_value_bytes(v)))
    return ko, vo


def exclusive_scan(data):
    d = np.array(data, dtype=np.uint32, copy=True)
    lib().orc_exclusive_scan_u32(_p(d), C.c_uint64(d.size))
    return d


def fill_leaves(nodes, ids):
    ids = np.ascontiguousarray(ids, dtype=np.uint32)
    lib().orc_fill_leaves(_p(nodes), _p(ids), C.c_uint32(len(ids)))


def generate_bvh(codes, nodes):
    codes = np.ascontiguousarray(codes, dtype=np.uint32)
    lib().orc_generate_bvh(_p(codes), _p(nodes), C.c_uint32(len(codes)))


def build_tree(codes_sorted, ids_sorted, fill=0):
    """leaves + internals; untouched fields (root parent, leaf data[1]) = fill."""
    n = len(codes_sorted)
    nodes = np.zeros(2 * n - 1, Node)
    nodes.view(np.uint32)[:] = fill
    fill_leaves(nodes, ids_sorted)
    generate_bvh(codes_sorted, nodes)
    return nodes


def leaf_bounds(bounds_arr, coords4, radii, nodes):
    n = (len(nodes) + 1) // 2
    getattr(lib(), "orc_leaf_bounds_" + _sfx(coords4.dtype))(
        _p(bounds_arr), _p(np.ascontiguousarray(coords4)), _p(np.ascontiguousarray(radii)), _p(nodes), C.c_uint32(n))


def internal_bounds(bounds_arr, nodes):
    n = (len(nodes) + 1) // 2
    flags = np.zeros(2 * n - 1, np.uint32)
    getattr(lib(), "orc_internal_bounds_" + _sfx(bounds_arr.dtype))(_p(bounds_arr), _p(flags), _p(nodes), C.c_uint32(n))


def refit(coords4, radii, nodes, fill=0.0):
    n = (len(nodes) + 1) // 2
    b = np.full((2 * n - 1, 2, 4), fill, coords4.dtype)
    leaf_bounds(b, coords4, radii, nodes)
    internal_bounds(b, nodes)
    return b


def traverse(nodes, bounds_arr, capacity=None, q_begin=0, q_end=None, want_counts=False):
    """Returns (total, pairs[min(total, capacity)], counts or None) in the
    sequential work-item order (query-major, DFS, child a before b)."""
    n = (len(nodes) + 1) // 2
    q_end = n if q_end is None else q_end
    fn = getattr(lib(), "orc_traverse_" + _sfx(bounds_arr.dtype))
    counts = np.zeros(n, np.uint32) if want_counts else None
    if capacity is None:
        capacity = int(fn(None, C.c_uint64(0), _p(nodes), _p(bounds_arr), C.c_uint32(n),
                          C.c_uint32(q_begin), C.c_uint32(q_end), None))
    pairs = np.zeros((max(capacity, 1), 2), np.uint32)
    total = int(fn(_p(pairs), C.c_uint64(capacity), _p(nodes), _p(bounds_arr), C.c_uint32(n),
                   C.c_uint32(q_begin), C.c_uint32(q_end), _p(counts)))
    return total, pairs[:min(total, capacity)], counts


def brute_force(coords4, radii):
    c = np.ascontiguousarray(coords4)
    r = np.ascontiguousarray(radii, dtype=c.dtype).reshape(-1)
    fn = getattr(lib(), "orc_brute_force_" + _sfx(c.dtype))
    total = int(fn(None, C.c_uint64(0), _p(c), _p(r), C.c_uint32(len(c))))
    pairs = np.zeros((max(total, 1), 2), np.uint32)
    fn(_p(pairs), C.c_uint64(total), _p(c), _p(r), C.c_uint32(len(c)))
    return pairs[:total]


def find_offsets(values, n_offsets, offset_dtype=np.uint32):
    v = np.ascontiguousarray(values)
    out = np.empty(n_offsets, offset_dtype)
    lib().orc_find_offsets(_p(v), C.c_uint64(len(v)), C.c_uint32(v.dtype.itemsize), _p(out),
                           C.c_uint64(n_offsets), C.c_uint32(out.dtype.itemsize))
    return out


def gather(in_values, indices, out_values):
    lib().orc_gather(_p(in_values), _p(indices), _p(out_values), C.c_uint64(len(indices)),
                     C.c_uint32(_value_bytes(in_values)), C.c_uint32(indices.dtype.itemsize))
    return out_values


def scatter(in_values, indices, out_values):
    lib().orc_scatter(_p(in_values), _p(indices), _p(out_values), C.c_uint64(len(indices)),
                      C.c_uint32(_value_bytes(in_values)), C.c_uint32(indices.dtype.itemsize))
    return out_values


# ---------------------------------------------------------------- pipeline
class Pipeline:
    """collision/collision.py:130-198 end to end on the host cores (float32).
    Holds its scratch so repeated timed calls do not re-allocate."""

    def __init__(self, n):
        self.n = n
        self.codes = np.empty(2 * n, np.uint32)
        self.ids = np.empty(2 * n, np.uint32)
        self.nodes = np.zeros(2 * n - 1, Node)
        self.bounds = np.zeros((2 * n - 1, 2, 4), np.float32)
        self.flags = np.zeros(2 * n - 1, np.uint32)
        self.offsets = np.zeros(n, np.uint64)

    def run(self, coords4, radii, capacity=0, single_walk=False, pairs=None):
        """Returns (total, pairs[:min(total, capacity)]).  single_walk=True: the reference's own
        protocol -- one walk per query, rows appended through a counter in ARBITRARY order
        (collision.cl:202-215); otherwise count -> scan -> write in the deterministic order.
        `pairs`: optional preallocated (capacity, 2) uint32 row buffer (timed loops)."""
        n = self.n
        assert coords4.shape == (n, 4) and coords4.dtype == np.float32
        radii = np.ascontiguousarray(radii, dtype=np.float32).reshape(-1)
        if pairs is None:
            pairs = np.zeros((max(capacity, 1), 2), np.uint32)
        assert pairs.dtype == np.uint32 and pairs.size >= 2 * capacity
        if single_walk:
            total = int(lib().orc_pipeline_f32_single(
                _p(coords4), _p(radii), C.c_uint32(n), _p(pairs), C.c_uint64(capacity),
                _p(self.codes), _p(self.ids), _p(self.nodes), _p(self.bounds), _p(self.flags)))
        else:
            total = int(lib().orc_pipeline_f32(
                _p(coords4), _p(radii), C.c_uint32(n), _p(pairs), C.c_uint64(capacity),
                _p(self.codes), _p(self.ids), _p(self.nodes), _p(self.bounds), _p(self.flags), _p(self.offsets)))
        return total, pairs[:min(total, capacity)]

    @property
    def sorted_codes(self):
        return self.codes[self.n:]

    @property
    def sorted_ids(self):
        return self.ids[self.n:]


def collide(coords4, radii, quirk_ngroups=None):
    """Stage-by-stage single pass through the oracle; returns a dict with every
    intermediate (range, codes, sorted codes/ids, nodes, bounds, total, pairs).
    quirk_ngroups=(ngroups, group_size) reproduces the reference's
    dropped-partial scene bounds for non-power-of-two ngroups."""
    c = np.ascontiguousarray(coords4)
    r = np.ascontiguousarray(radii, dtype=c.dtype).reshape(-1)
    n = len(c)
    ng, gs = quirk_ngroups if quirk_ngroups else (0, 0)
    rng = bounds(c, ng, gs)
    codes = morton(c, rng)
    scodes, sids = sort_pairs(codes, np.arange(n, dtype=np.uint32))
    nodes = build_tree(scodes, sids)
    b = refit(c, r, nodes)
    total, pairs, counts = traverse(nodes, b, want_counts=True)
    return dict(range=rng, codes=codes, sorted_codes=scodes, sorted_ids=sids, nodes=nodes,
                bounds=b, total=total, pairs=pairs, counts=counts)


# -------------------------------------------------------------------- _ref
_REF = {}


def ref(dtype=np.float32):
    """oracle/_ref/libref_{f32,f64}.so, or None when it was not built."""
    key = _sfx(dtype)
    if key not in _REF:
        so = _HERE / "_ref" / "libref_{}.so".format(key)
        _REF[key] = C.CDLL(str(so)) if so.exists() else None
    return _REF[key]
