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
import ctypes as C

from numpy import dtype

from . import _lib
from .misc import Program, dtype_lanes, dtype_sizeof, same_context
from .misc import np_float_dtypes as _float_names

np_float_dtypes = set(map(dtype, _float_names))


_ACC_FUNCS = {"min": 0, "fmin": 0, "max": 1, "fmax": 1, "ADD": 2}
_ALL_BASES = ('float32', 'float64', 'uint32', 'int32', 'uint64', 'int64')


def _parse_initialiser(text):
    """The OpenCL-C initialiser strings upstream puts in accumulator lists ("INFINITY",
    "-INFINITY", "0", numeric literals) as a float."""
    t = str(text).strip().replace(" ", "")
    table = {"INFINITY": float("inf"), "+INFINITY": float("inf"), "-INFINITY": float("-inf"),
             "NAN": float("nan"), "FLT_MAX": 3.4028234663852886e38, "-FLT_MAX": -3.4028234663852886e38,
             "DBL_MAX": 1.7976931348623157e308, "-DBL_MAX": -1.7976931348623157e308}
    if t in table:
        return table[t]
    try:
        return float(t.rstrip("fFuUlL") or t)
    except ValueError:
        raise ValueError("Unsupported accumulator initialiser: {!r}".format(text))


class ReductionProgram(Program):
    """collision/reduce.py:9-22.  A subclass names its `accumulator` list of (initialiser,
    function) pairs exactly as upstream; functions: min / max (fmin / fmax) and ADD.  The two
    lists the package ships (Bounds, Summer) also name a single-launch entry point (`_entry`);
    any other list runs through the generic bounds1 / bounds2 pair with the reference's
    ngroups x group_size geometry (csrc/compat.cu)."""
    accumulator = []          # [(initialiser, function)], as upstream
    _entry = None             # single-launch C-ABI entry point of a shipped list (else: generic pair)
    _value_bases = ('float32', 'float64')   # scalar types `_entry` is instantiated for

    def __init__(self, ctx, value_dtype):
        self.value_dtype = dtype(value_dtype)
        self.acc_dtype = dtype((self.value_dtype, len(self.accumulator)))
        if not self.accumulator:
            raise ValueError("ReductionProgram needs a non-empty accumulator list")
        if len(self.accumulator) > 8:
            raise ValueError("At most 8 accumulators are supported")
        bases = self._value_bases if self._entry else _ALL_BASES
        if self.value_dtype.base.name not in bases:
            raise ValueError("Invalid dtype: {}".format(self.value_dtype))
        dtype_sizeof(self.value_dtype)   # validates the vector width
        inits, funcs = zip(*self.accumulator)
        for f in funcs:
            if f not in _ACC_FUNCS:
                raise ValueError("Unsupported accumulator function: {!r} (min, max, ADD)".format(f))
        self._ops = (C.c_int * len(funcs))(*[_ACC_FUNCS[f] for f in funcs])
        self._inits = (C.c_double * len(inits))(*[_parse_initialiser(i) for i in inits])
        super().__init__(ctx)
        L, P = _lib.lib(), _lib.ptr
        code, lanes, n_acc = _lib.value_code(self.value_dtype), dtype_lanes(self.value_dtype), len(funcs)

        # kernel-level contracts of collision/reduce.cl:5-58 (collision/reduce.py:10-11)
        def bounds1(cq, gs, ls, values, n, group_accs, _scratch):
            _lib.check(L.col_compat_reduce1(P(values), int(n), code, lanes, self._ops, self._inits, n_acc,
                                            P(group_accs), gs // ls, ls, _lib.stream_ptr(cq)), "col_compat_reduce1")

        def bounds2(cq, gs, ls, group_accs, output):
            _lib.check(L.col_compat_reduce2(P(group_accs), P(output), code, lanes, self._ops, n_acc, gs,
                                            _lib.stream_ptr(cq)), "col_compat_reduce2")

        self.kernels = {'bounds1': _lib.kernel_launcher(bounds1), 'bounds2': _lib.kernel_launcher(bounds2)}


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
        self._group_buf = None  # per-group partials of the generic two-kernel path

    def _ensure_scratch(self, cq=None):
        if self._scratch is None:
            self._scratch = _lib.alloc_on(cq, self.program.context, _lib.lib().col_reduce_scratch_bytes())
        return self._scratch

    def resize(self, ngroups=None, group_size=None):
        if ngroups is not None:
            self.ngroups = ngroups
        if group_size is not None:
            self.group_size = group_size

    def reduce(self, cq, size, values_buf, output_buf, wait_for=None):
        if self.program._entry is None:
            return self._reduce_generic(cq, size, values_buf, output_buf, wait_for)
        _lib.begin(cq, wait_for)
        fn = getattr(_lib.lib(), self.program._entry)
        _lib.check(fn(_lib.ptr(values_buf), size, _lib.value_code(self.program.value_dtype),
                      dtype_lanes(self.program.value_dtype), _lib.ptr(output_buf),
                      _lib.ptr(self._ensure_scratch(cq)), _lib.stream_ptr(cq)), self.program._entry)
        return _lib.done(cq)

    def _reduce_generic(self, cq, size, values_buf, output_buf, wait_for):
        """Any accumulator list: upstream's two launches with upstream's geometry
        (collision/reduce.py:62-76) -- bounds1 over ngroups x group_size work-items,
        bounds2 over one group of ngroups."""
        need = self.ngroups * dtype_sizeof(self.program.acc_dtype)
        if self._group_buf is None or self._group_buf.numel() < need:
            self._group_buf = _lib.alloc_on(cq, self.program.context, need)
        e = self.program.kernels['bounds1'](cq, (self.ngroups,), (self.group_size,), values_buf, size,
                                            self._group_buf, None, g_times_l=True, wait_for=wait_for)
        return self.program.kernels['bounds2'](cq, (1,), (self.ngroups,), self._group_buf, output_buf,
                                               g_times_l=True, wait_for=[e])
