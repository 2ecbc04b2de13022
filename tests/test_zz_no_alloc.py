"""col_exclusive_scan_u32_static -- the scan with its scratch in a device array built into the
library (no allocation by the callee) -- against col_exclusive_scan_u32 on the same inputs,
through the C ABI.  Bit-exact (u32 arithmetic mod 2^32, u64 grand total).  Collected last on
purpose (the file name): the array is shared by every call in the process, so this also runs
after whatever the other tests left on the GPU."""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu

MAX_N = 1 << 24      # COL_SCAN_STATIC_MAX_N in include/collision_b200.h


def _both(L, P, dev, cq, src):
    """(static result, static total, scratch result, scratch total) for one input."""
    import torch
    from collision_b200 import _lib
    n = src.numel()
    a, b = src.clone(), src.clone()
    ta = torch.full((1,), -1, dtype=torch.int64, device=dev)
    tb = torch.full((1,), -1, dtype=torch.int64, device=dev)
    scratch = _lib.alloc_bytes(dev, L.col_scan_scratch_bytes(n))
    cq.wait_stream(torch.cuda.current_stream(dev))
    assert L.col_exclusive_scan_u32_static(P(a), n, P(ta), cq.cuda_stream) == 0
    assert L.col_exclusive_scan_u32(P(b), n, P(tb), P(scratch), cq.cuda_stream) == 0
    cq.synchronize()
    return a, int(ta.item()), b, int(tb.item())


def test_static_scan_matches_scratch_scan_twice(cuda_env):
    import torch
    from collision_b200 import _lib
    dev, cq = cuda_env
    L, P = _lib.lib(), _lib.ptr
    g = torch.Generator(device=dev).manual_seed(20)
    # two calls in a row on the shared array: the second is SHORTER and of other data, so flags or
    # aggregates left over from the first would show; then ragged, one-tile and full-size cases
    for n in ((1 << 22) + 4097, (1 << 20) + 13, 8192, 5, MAX_N):
        wide = torch.randint(0, 1 << 32, (n,), dtype=torch.int64, device=dev, generator=g)
        src = wide.to(torch.int32)
        a, ta, b, tb = _both(L, P, dev, cq, src)
        assert torch.equal(a, b), n
        assert ta == tb == int(wide.sum().item()), n
        want = torch.cumsum(wide, 0)
        assert a[0].item() == 0
        assert torch.equal(a[1:].to(torch.int64) & 0xFFFFFFFF, want[:-1] & 0xFFFFFFFF), n


def test_static_scan_same_input_twice(cuda_env):
    """The array keeps the previous call's contents: the same call repeated gives the same answer."""
    import torch
    from collision_b200 import _lib
    dev, cq = cuda_env
    L, P = _lib.lib(), _lib.ptr
    n = 3 * 8192 + 1
    src = torch.arange(n, dtype=torch.int32, device=dev) % 1000
    first = _both(L, P, dev, cq, src)
    second = _both(L, P, dev, cq, src)
    assert torch.equal(first[0], first[2]) and torch.equal(second[0], second[2])
    assert torch.equal(first[0], second[0]) and first[1] == second[1] == first[3]


def test_static_scan_limits(cuda_env):
    import torch
    from collision_b200 import _lib
    dev, cq = cuda_env
    L, P = _lib.lib(), _lib.ptr
    buf = torch.zeros(16, dtype=torch.int32, device=dev)
    tot = torch.full((1,), 7, dtype=torch.int64, device=dev)
    cq.wait_stream(torch.cuda.current_stream(dev))
    # beyond the built-in array: refused before anything is launched, buffer untouched
    assert L.col_exclusive_scan_u32_static(P(buf), MAX_N + 1, P(tot), cq.cuda_stream) == -2
    assert L.col_exclusive_scan_u32_static(None, 4, P(tot), cq.cuda_stream) == -1
    cq.synchronize()
    assert int(tot.item()) == 7 and int(buf.abs().sum().item()) == 0
    # n = 0: only the total is written
    assert L.col_exclusive_scan_u32_static(None, 0, P(tot), cq.cuda_stream) == 0
    cq.synchronize()
    assert int(tot.item()) == 0
    with pytest.raises(_lib.CollisionLibraryError, match="too large"):
        _lib.check(L.col_exclusive_scan_u32_static(P(buf), C.c_uint64(1 << 40), None, cq.cuda_stream), "static scan")
