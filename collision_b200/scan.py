"""Exclusive prefix sum of uint32 (mirror of collision/scan.py upstream).

PrefixScanner keeps the reference's constructor, validation, `block_lengths`
property and `prefix_sum(cq, values_buf, wait_for)` contract (in place,
exclusive, wraps mod 2^32), but one sm_100a decoupled-look-back kernel
(csrc/scan.cu) replaces the 2*levels+1 launches of scan.py:75-112.
`group_size` is validated exactly as upstream and otherwise only a hint.
"""
from numpy import dtype

from . import _lib
from .misc import Program, nextPowerOf2, roundUp, same_context


def ceildiv(a, b):
    return -(-a // b)


class PrefixScanProgram(Program):
    """collision/scan.py:10-13: a handle (the kernel is prebuilt)."""

    def __init__(self, ctx):
        super().__init__(ctx)
        L, P = _lib.lib(), _lib.ptr

        # kernel-level contracts of collision/scan.cl:5-36 (tests/test_scan.py:8-21 upstream)
        def local_scan(cq, gs, ls, data, _local, block_sums):
            _lib.check(L.col_compat_local_scan(P(data), P(block_sums), gs // ls, 2 * ls, _lib.stream_ptr(cq)),
                       "col_compat_local_scan")

        def block_scan(cq, gs, ls, data, block_sums):
            _lib.check(L.col_compat_block_scan(P(data), P(block_sums), 2 * gs, 2 * ls, _lib.stream_ptr(cq)),
                       "col_compat_block_scan")

        self.kernels = {'local_scan': _lib.kernel_launcher(local_scan), 'block_scan': _lib.kernel_launcher(block_scan)}


class PrefixScanner:
    block_sums_dtype = dtype('uint32')

    def __init__(self, ctx, size, group_size, program=None):
        self.check_size(size, group_size)
        self.size = size
        self.group_size = group_size
        if program is None:
            program = PrefixScanProgram(ctx)
        elif not same_context(program.context, ctx):
            raise ValueError("Scanner and program context must match")
        self.program = program
        self._scratch = None    # device scratch is allocated on first use

    def _ensure_scratch(self, cq=None):
        need = _lib.lib().col_scan_scratch_bytes(self.size)
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = _lib.alloc_on(cq, self.program.context, need)
        return self._scratch

    @staticmethod
    def check_size(size, group_size):
        # collision/scan.py:34-39
        if group_size != nextPowerOf2(group_size):
            raise ValueError("Group size ({}) must be a power of two".format(group_size))
        if size % (group_size * 2):
            raise ValueError("Size must be multiple of 2 * group_size ({})".format(group_size))

    def resize(self, size=None, group_size=None):
        size = self.size if size is None else size
        group_size = self.group_size if group_size is None else group_size
        self.check_size(size, group_size)
        self.size, self.group_size = size, group_size

    @property
    def block_lengths(self):
        """Level plan of the reference's recursive scan (collision/scan.py:62-73).
        Kept as a derived property (tests/test_scan_py.py:32-45 upstream); the
        single-pass kernel here has no levels."""
        two_g = self.group_size * 2
        lengths = []
        size = roundUp(ceildiv(self.size, two_g), two_g)
        while size > two_g:
            size = roundUp(size, two_g)
            lengths.append(size)
            size = ceildiv(size, two_g)
        lengths.append(nextPowerOf2(size))
        return tuple(lengths)

    def prefix_sum(self, cq, values_buf, wait_for=None):
        _lib.begin(cq, wait_for)
        _lib.check(_lib.lib().col_exclusive_scan_u32(
            _lib.ptr(values_buf), self.size, None, _lib.ptr(self._ensure_scratch(cq)), _lib.stream_ptr(cq)),
            "col_exclusive_scan_u32")
        return _lib.done(cq)
