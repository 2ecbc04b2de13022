"""ctypes binding of libcollision_b200.so (the C ABI of include/collision_b200.h).

There is NO fallback path: if the library is missing or a call fails, this
module raises.  torch is used for device memory, streams and events only.
"""
import ctypes as C
from pathlib import Path

import numpy as np
import torch

_HERE = Path(__file__).resolve().parent
import os as _os
# COL_B200_DEBUG=1: the -DCOL_DEBUG build (python -m collision_b200.build --debug)
_LIB_PATH = _HERE / ("libcollision_b200_debug.so" if _os.environ.get("COL_B200_DEBUG") == "1" else "libcollision_b200.so")
_lib = None

COL_F32, COL_F64 = 0, 1
COL_U32, COL_I32, COL_U64, COL_I64 = 2, 3, 4, 5

_vp, _u64, _u32, _int, _sz = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int, C.c_size_t

# name -> (restype, argtypes); mirrors include/collision_b200.h one to one
SIGNATURES = {
    "col_abi_version": (_int, []),
    "col_built_arch": (_int, []),
    "col_reduce_scratch_bytes": (_sz, []),
    "col_bounds": (_int, [_vp, _u64, _int, _u32, _vp, _vp, _vp]),
    "col_sum": (_int, [_vp, _u64, _int, _u32, _vp, _vp, _vp]),
    "col_morton": (_int, [_vp, _vp, _vp, _vp, _int, _u32, _u32, _vp]),
    "col_scan_scratch_bytes": (_sz, [_u64]),
    "col_exclusive_scan_u32": (_int, [_vp, _u64, _vp, _vp, _vp]),
    "col_exclusive_scan_u32_static": (_int, [_vp, _u64, _vp, _vp]),
    "col_sort_scratch_bytes": (_sz, [_u64, _u32, _u32]),
    "col_sort_pairs": (_int, [_vp, _vp, _vp, _vp, _u64, _u32, _u32, _u32, _vp, _vp]),
    "col_radix_pass": (_int, [_vp, _vp, _vp, _vp, _u64, _u32, _u32, _u32, _u32, _vp, _vp]),
    "col_build_tree": (_int, [_vp, _vp, _vp, _u32, _vp]),
    "col_fill_leaves": (_int, [_vp, _vp, _u32, _vp]),
    "col_refit": (_int, [_vp, _vp, _vp, _vp, _vp, _int, _u32, _vp]),
    "col_leaf_bounds": (_int, [_vp, _vp, _vp, _vp, _int, _u32, _vp]),
    "col_traverse_count": (_int, [_vp, _vp, _vp, _int, _u32, _u32, _u32, _vp]),
    "col_traverse_write": (_int, [_vp, _u64, _vp, _u64, _vp, _vp, _int, _u32, _u32, _u32, _vp]),
    "col_traverse_atomic": (_int, [_vp, _vp, _u32, _vp, _vp, _int, _u32, _u32, _u32, _vp]),
    "col_build_tree_wide": (_int, [_vp, _vp, _vp, _vp, _vp, _u32, _vp]),
    "col_refit_wide": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _int, _u32, _vp]),
    "col_build_refit_wide": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _int, _u32, _vp]),
    "col_traverse_scratch_bytes": (_sz, [_u32, _u64]),
    "col_traverse_wide_count": (_int, [_vp, _vp, _vp, _vp, _vp, _int, _u32, _u32, _u32, _vp]),
    "col_traverse_wide_pairs": (_int, [_vp, _u64, _u64, _vp, _vp, _vp, _vp, _vp, _int, _u32, _u32, _u32, _vp, _vp]),
    "col_traverse_wide_pairs_marked": (_int, [_vp, _u64, _u64, _vp, _vp, _vp, _vp, _vp, _int, _u32, _u32, _u32, _vp, _vp, _vp]),
    "col_traverse_wide_write": (_int, [_vp, _u64, _vp, _u64, _vp, _vp, _vp, _vp, _int, _u32, _u32, _u32, _vp]),
    "col_find_offsets": (_int, [_vp, _u64, _u32, _vp, _u64, _u32, _vp]),
    "col_gather": (_int, [_vp, _vp, _vp, _u64, _u32, _u32, _vp]),
    "col_scatter": (_int, [_vp, _vp, _vp, _u64, _u32, _u32, _vp]),
    "col_compat_local_scan": (_int, [_vp, _vp, _u64, _u32, _vp]),
    "col_compat_block_scan": (_int, [_vp, _vp, _u64, _u32, _vp]),
    "col_compat_block_sort": (_int, [_vp, _vp, _vp, _u64, _u32, _u32, _u32, _u32, _u32, _vp]),
    "col_compat_scatter": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _u64, _u32, _u32, _u32, _u32, _u32, _vp]),
    "col_compat_find_offsets": (_int, [_vp, _u64, _u32, _vp, _u32, _vp]),
    "col_compat_reduce1": (_int, [_vp, _u64, _int, _u32, _vp, _vp, _u32, _vp, _u32, _u32, _vp]),
    "col_compat_reduce2": (_int, [_vp, _vp, _int, _u32, _vp, _u32, _u32, _vp]),
    "col_ipc_alloc": (_int, [_vp, _u64]),
    "col_ipc_free": (_int, [_vp]),
    "col_ipc_export": (_int, [_vp, _vp]),
    "col_ipc_open": (_int, [_vp, _vp]),
    "col_ipc_close": (_int, [_vp]),
    "col_shard_packed_bytes": (_sz, [_u32, _int]),
    "col_shard_list_bytes": (_sz, [_u32]),
    "col_shard_build_windows": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _int, _u32, _u32, _u32, _vp, _u32, _vp]),
    "col_shard_exchange_bytes": (_sz, [_u32, _int]),
    "col_shard_top_local": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _int, _u32, _u32, _u32, _vp, _u32, _vp, _u32, _vp]),
    "col_shard_top": (_int, [_vp, _u32, _u32, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _int, _u32, _vp, _vp]),
    "col_shard_halo": (_int, [_vp, _u32, _vp, _u32, _u32, _u32, _vp, _vp, _vp, _int, _vp, _vp, _u32, _vp, _vp, _vp, _u32, _u32,
                              _u32, _vp, _vp, _vp]),
    "col_shard_sort_scratch_bytes": (_sz, [_u64]),
    "col_shard_sort": (_int, [_vp, _u32, _u32, _u32, _u32, _vp, _vp, _vp, _vp, _vp]),
    "col_shard_count": (_int, [_vp, _u32, _u32, _vp, _vp]),
    "col_shard_push_scratch_bytes": (_sz, [_u64, _u32]),
    "col_shard_push": (_int, [_vp, _u32, _u32, _vp, _u32, _vp, _u32, _u32, _vp, _vp, _int, _u32, _vp, _vp, _vp, _vp, _vp,
                              _vp, _vp]),
    "col_shard_sort_received_scratch_bytes": (_sz, [_u64]),
    "col_shard_sort_received": (_int, [_vp, _vp, _u32, _vp, _u32, _vp, _vp, _vp, _vp, _vp]),
    "col_fill_u32": (_int, [_vp, _u32, _u64, _vp]),
    "col_range": (_int, [_vp, _u64, _vp]),
    "col_copy": (_int, [_vp, _vp, _u64, _vp]),
}


class CollisionLibraryError(RuntimeError):
    pass


def lib_path():
    return _LIB_PATH


def lib():
    """The loaded library.  Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not _LIB_PATH.exists():
            raise CollisionLibraryError(
                "{} not found: build it with `python -m collision_b200.build` "
                "(nvcc, sm_100a).  collision_b200 has no CPU or PyTorch fallback.".format(_LIB_PATH))
        handle = C.CDLL(str(_LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)   # AttributeError if the ABI is incomplete
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc, what):
    if rc == 0:
        return
    if rc > 0:
        raise CollisionLibraryError("{}: CUDA error {} ({})".format(what, rc, _cuda_error_name(rc)))
    raise CollisionLibraryError("{}: {}".format(what, {-1: "bad argument", -2: "size too large",
                                                     -3: "unsupported dtype"}.get(rc, rc)))


def _cuda_error_name(rc):
    try:
        rt = C.CDLL("libcudart.so")
        rt.cudaGetErrorName.restype = C.c_char_p
        return rt.cudaGetErrorName(C.c_int(rc)).decode()
    except Exception:
        return "?"


# ------------------------------------------------------------ torch helpers
def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return None
    if not isinstance(t, torch.Tensor):
        raise TypeError("expected a torch.Tensor device buffer, got {}".format(type(t)))
    if not t.is_cuda:
        raise ValueError("buffer must live on a CUDA device")
    if not t.is_contiguous():
        raise ValueError("buffer must be contiguous")
    return t.data_ptr()


def stream_ptr(cq):
    return cq.cuda_stream


def begin(cq, wait_for):
    """Make `cq` wait for the given events (the reference's wait_for lists)."""
    for e in wait_for or ():
        if e is not None:
            cq.wait_event(e)


def done(cq):
    """Event marking everything enqueued on `cq` so far (the reference returns cl.Event)."""
    ev = torch.cuda.Event()
    ev.record(cq)
    return ev


def kernel_launcher(fn):
    """A `program.kernels[name]` entry with PyOpenCL's call convention
    kernel(cq, global_size, local_size, *args, g_times_l=False, wait_for=None) -> event.
    `fn(cq, n_threads, local_size_or_None, *args)` enqueues the work; cl.LocalMemory
    arguments of the reference's kernels are accepted as any placeholder and ignored."""
    def launch(cq, global_size, local_size, *args, g_times_l=False, wait_for=None, global_offset=None):
        gs = int(global_size[0]) if not isinstance(global_size, int) else global_size
        ls = None if local_size is None else (int(local_size[0]) if not isinstance(local_size, int) else local_size)
        if g_times_l and ls:
            gs *= ls
        begin(cq, wait_for)
        fn(cq, gs, ls, *args)
        return done(cq)
    return launch


def device_of(ctx):
    dev = torch.device(ctx)
    if dev.type != "cuda":
        raise ValueError("collision_b200 runs on CUDA devices only (got {})".format(dev))
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    return dev


_DEBUG = _os.environ.get("COL_B200_DEBUG") == "1"
_GUARD = 4096            # bytes of canary before and after every scratch buffer of a debug run
_guarded = []            # weak references to the guarded allocations


def alloc_bytes(ctx, nbytes):
    """Private scratch (the reference's HOST_NO_ACCESS cl.Buffer).  With COL_B200_DEBUG=1 every
    buffer sits between two 4 KB canaries that check_canaries() verifies (compute-sanitizer is
    not available on every pool; this catches out-of-bounds writes of the kernels' scratch)."""
    n = max(int(nbytes), 1)
    if not _DEBUG:
        return torch.empty(n, dtype=torch.uint8, device=ctx)
    import weakref
    whole = torch.empty(n + 2 * _GUARD, dtype=torch.uint8, device=ctx)
    whole[:_GUARD].fill_(0xA5)
    whole[_GUARD + n:].fill_(0xA5)
    _guarded.append(weakref.ref(whole))
    view = whole[_GUARD:_GUARD + n]
    view._col_whole = whole        # keeps the canaries alive as long as the view is
    return view


def check_canaries():
    """Debug runs: every live scratch buffer's canaries are intact.  Returns how many were checked."""
    torch.cuda.synchronize()
    live = 0
    for ref in list(_guarded):
        whole = ref()
        if whole is None:
            _guarded.remove(ref)
            continue
        live += 1
        if not (bool((whole[:_GUARD] == 0xA5).all()) and bool((whole[-_GUARD:] == 0xA5).all())):
            raise CollisionLibraryError("scratch canary overwritten: a kernel wrote outside its buffer "
                                        "({} bytes allocation)".format(whole.numel() - 2 * _GUARD))
    return live


def alloc_on(cq, ctx, nbytes):
    """Scratch used only by work on stream `cq`: allocated with `cq` current, so the caching
    allocator hands a dropped block out again in cq's order (None: the current stream)."""
    if cq is None:
        return alloc_bytes(ctx, nbytes)
    with torch.cuda.stream(cq):
        return alloc_bytes(ctx, nbytes)


def float_code(dt):
    dt = np.dtype(dt).base
    if dt == np.float32:
        return COL_F32
    if dt == np.float64:
        return COL_F64
    raise ValueError("Invalid dtype: {}".format(dt))


_VALUE_CODES = {"float32": COL_F32, "float64": COL_F64, "uint32": COL_U32, "int32": COL_I32,
                "uint64": COL_U64, "int64": COL_I64}


def value_code(dt):
    """Value-type code of the reductions (floats for min/max, floats and integers for sums)."""
    name = np.dtype(dt).base.name
    if name not in _VALUE_CODES:
        raise ValueError("Invalid dtype: {}".format(name))
    return _VALUE_CODES[name]
