"""First index of each value in a sorted array (mirror of collision/offset.py upstream)."""
from numpy import dtype

from . import _lib
from .misc import Program, same_context
from .misc import np_unsigned_dtypes as _unsigned_names

np_unsigned_dtypes = set(map(dtype, _unsigned_names))


class OffsetProgram(Program):
    def __init__(self, ctx, value_dtype=dtype('uint32'), offset_dtype=dtype('uint32')):
        self.value_dtype = dtype(value_dtype)
        self.offset_dtype = dtype(offset_dtype)
        if self.value_dtype not in np_unsigned_dtypes:
            raise ValueError("Invalid value dtype: {}".format(self.value_dtype))
        if self.offset_dtype not in np_unsigned_dtypes:
            raise ValueError("Invalid offset dtype: {}".format(self.offset_dtype))
        for dt in (self.value_dtype, self.offset_dtype):
            if dt.itemsize not in (4, 8):
                raise ValueError("Unsupported dtype: {} (uint32 / uint64)".format(dt))
        super().__init__(ctx)
        L, P = _lib.lib(), _lib.ptr
        vb, ob = self.value_dtype.itemsize, self.offset_dtype.itemsize

        # kernel-level contract of collision/offset.cl:3-12 (tests/test_offset.py:12-47 upstream):
        # n_values - 1 work-items, no fill
        def find_offsets(cq, gs, ls, values, offsets):
            _lib.check(L.col_compat_find_offsets(P(values), gs, vb, P(offsets), ob, _lib.stream_ptr(cq)),
                       "col_compat_find_offsets")

        self.kernels = {'find_offsets': _lib.kernel_launcher(find_offsets)}


class OffsetFinder:
    def __init__(self, ctx, value_dtype=dtype('uint32'), offset_dtype=dtype('uint32'), program=None):
        if program is None:
            program = OffsetProgram(ctx, value_dtype, offset_dtype)
        else:
            if not same_context(program.context, ctx):
                raise ValueError("Sorter and program contexts must match")
            if program.value_dtype != value_dtype:
                raise ValueError("Sorter and program value dtypes must match")
            if program.offset_dtype != offset_dtype:
                raise ValueError("Sorter and program offset dtypes must match")
        self.program = program

    def find_offsets(self, cq, values_buf, n_values, offsets_buf, n_offsets, wait_for=None):
        """offsets[v] = first i with values[i] >= v; n_values above the maximum
        (collision/offset.py:40-49: host fill + find_offsets kernel, here one launch)."""
        _lib.begin(cq, wait_for)
        _lib.check(_lib.lib().col_find_offsets(
            _lib.ptr(values_buf), n_values, self.program.value_dtype.itemsize,
            _lib.ptr(offsets_buf), n_offsets, self.program.offset_dtype.itemsize,
            _lib.stream_ptr(cq)), "col_find_offsets")
        return _lib.done(cq)
