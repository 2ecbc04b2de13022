"""Indexed copy (mirror of collision/index.py upstream): gather / scatter."""
from numpy import dtype

from . import _lib
from .misc import Program, dtype_sizeof, same_context
from .misc import np_unsigned_dtypes as _unsigned_names

np_unsigned_dtypes = set(map(dtype, _unsigned_names))


class IndexProgram(Program):
    def __init__(self, ctx, value_dtype=dtype('uint32'), index_dtype=dtype('uint32')):
        self.value_dtype = dtype(value_dtype)
        self.index_dtype = dtype(index_dtype)
        if self.index_dtype not in np_unsigned_dtypes:
            raise ValueError("Invalid index dtype: {}".format(self.index_dtype))
        if self.index_dtype.itemsize not in (4, 8):
            raise ValueError("Unsupported index dtype: {} (uint32 / uint64)".format(self.index_dtype))
        super().__init__(ctx)
        L, P = _lib.lib(), _lib.ptr
        vb, ib = dtype_sizeof(self.value_dtype), self.index_dtype.itemsize

        # kernel-level contracts of collision/index.cl:1-13: one work-item per element
        def gather(cq, gs, ls, in_values, indices, out_values):
            _lib.check(L.col_gather(P(in_values), P(indices), P(out_values), gs, vb, ib, _lib.stream_ptr(cq)), "col_gather")

        def scatter(cq, gs, ls, in_values, indices, out_values):
            _lib.check(L.col_scatter(P(in_values), P(indices), P(out_values), gs, vb, ib, _lib.stream_ptr(cq)), "col_scatter")

        self.kernels = {'gather': _lib.kernel_launcher(gather), 'scatter': _lib.kernel_launcher(scatter)}


class Indexer:
    def __init__(self, ctx, value_dtype=dtype('uint32'), index_dtype=dtype('uint32'), program=None):
        if program is None:
            program = IndexProgram(ctx, value_dtype, index_dtype)
        else:
            if not same_context(program.context, ctx):
                raise ValueError("Sorter and program contexts must match")
            if program.index_dtype != index_dtype:
                raise ValueError("Sorter and program index dtypes must match")
            if program.value_dtype != value_dtype:
                raise ValueError("Sorter and program value dtypes must match")
        self.program = program

    def _run(self, name, cq, size, in_values_buf, indices_buf, out_values_buf, wait_for):
        _lib.begin(cq, wait_for)
        fn = getattr(_lib.lib(), name)
        _lib.check(fn(_lib.ptr(in_values_buf), _lib.ptr(indices_buf), _lib.ptr(out_values_buf), size,
                      dtype_sizeof(self.program.value_dtype), self.program.index_dtype.itemsize,
                      _lib.stream_ptr(cq)), name)
        return _lib.done(cq)

    def gather(self, cq, size, in_values_buf, indices_buf, out_values_buf, wait_for=None):
        """out[i] = in[indices[i]]   (collision/index.cl:1-6)"""
        return self._run("col_gather", cq, size, in_values_buf, indices_buf, out_values_buf, wait_for)

    def scatter(self, cq, size, in_values_buf, indices_buf, out_values_buf, wait_for=None):
        """out[indices[i]] = in[i]   (collision/index.cl:8-13)"""
        return self._run("col_scatter", cq, size, in_values_buf, indices_buf, out_values_buf, wait_for)
