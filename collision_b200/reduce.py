"""Generic vector reduction (mirror of collision/reduce.py upstream).

Upstream renders reduce.cl through jinja2 from an accumulator list and runs two
kernels (bounds1 over ngroups x group_size threads, bounds2 over the partials:
reduce.py:62-76).  Here the two accumulator lists the package ships -- min/max
(Bounds) and sum (Summer) -- are template instantiations of one single-launch
kernel (csrc/reduce.cu).  `ngroups` / `group_size` are accepted for API parity
and do not shape the launch; consequently the true min/max is returned for any
ngroups (upstream drops tail partials when ngroups is not a power of two,
reduce.cl:42-52 -- see DESIGN.md).
"""
from numpy import dtype

from . import _lib
from .misc import Program, dtype_lanes, dtype_sizeof, same_context
from .misc import np_float_dtypes as _float_names

np_float_dtypes = set(map(dtype, _float_names))


class ReductionProgram(Program):
    accumulator = []          # [(initialiser, function)], as upstream
    _entry = None             # name of the C-ABI entry point
    _value_bases = ('float32', 'float64')   # scalar types the entry point is instantiated for

    def __init__(self, ctx, value_dtype):
        self.value_dtype = dtype(value_dtype)
        self.acc_dtype = dtype((self.value_dtype, len(self.accumulator)))
        if self.value_dtype.base.name not in self._value_bases:
            raise ValueError("Invalid dtype: {}".format(self.value_dtype))
        dtype_sizeof(self.value_dtype)   # validates the vector width
        super().__init__(ctx)


class Reducer:
    program_type = ReductionProgram

    def __init__(self, ctx, ngroups, group_size, value_dtype, program=None):
        if program is None:
            program = self.program_type(ctx, value_dtype)
        else:
            if not same_context(program.context, ctx):
                raise ValueError("Collider and program context must match")
            if program.value_dtype != value_dtype:
                raise ValueError("Reducer and program value dtypes must match")
        self.program = program
        self.ngroups = ngroups
        self.group_size = group_size
        self._scratch = None    # device scratch is allocated on first use

    def _ensure_scratch(self):
        if self._scratch is None:
            self._scratch = _lib.alloc_bytes(self.program.context, _lib.lib().col_reduce_scratch_bytes())
        return self._scratch

    def resize(self, ngroups=None, group_size=None):
        if ngroups is not None:
            self.ngroups = ngroups
        if group_size is not None:
            self.group_size = group_size

    def reduce(self, cq, size, values_buf, output_buf, wait_for=None):
        _lib.begin(cq, wait_for)
        fn = getattr(_lib.lib(), self.program._entry)
        _lib.check(fn(_lib.ptr(values_buf), size, _lib.value_code(self.program.value_dtype),
                      dtype_lanes(self.program.value_dtype), _lib.ptr(output_buf),
                      _lib.ptr(self._ensure_scratch()), _lib.stream_ptr(cq)), self.program._entry)
        return _lib.done(cq)
