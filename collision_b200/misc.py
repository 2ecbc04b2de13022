"""Host-side sizing and dtype helpers (mirror of collision/misc.py:28-71 upstream).

The reference's Program / SimpleProgram (misc.py:6-25) JIT-compile OpenCL text;
here a Program is a light handle: the kernels live in the prebuilt sm_100a
library, the handle only carries the context (a CUDA device) and the dtypes a
program was "built" for, so that sharing programs between objects and the
mismatch ValueErrors keep working.
"""
import operator as op
from functools import reduce

from numpy import dtype

from . import _lib


class Program:
    kernel_args = {}

    def __init__(self, ctx, options=None):
        self._context = _lib.device_of(ctx)
        self.options = list(options or [])
        _lib.lib()   # fail now, loudly, if the CUDA library is not there
        self.kernels = {}

    @property
    def context(self):
        return self._context


SimpleProgram = Program


def same_context(a, b):
    return _lib.device_of(a) == _lib.device_of(b)


def roundUp(x, base=1):
    return -(-x // base) * base


def nextPowerOf2(x):
    return 1 << (x - 1).bit_length()


def product(xs):
    return reduce(op.mul, xs, 1)


np_integer_dtypes = ['int8', 'int16', 'int32', 'int64']
np_unsigned_dtypes = ['uint8', 'uint16', 'uint32', 'uint64']
np_float_dtypes = ['float16', 'float32', 'float64']
np_dtypes = np_integer_dtypes + np_unsigned_dtypes + np_float_dtypes

c_integer_dtypes = ['char', 'short', 'int', 'long']
c_unsigned_dtypes = ['u' + t for t in c_integer_dtypes]
c_float_dtypes = ['half', 'float', 'double']
c_dtypes = c_integer_dtypes + c_unsigned_dtypes + c_float_dtypes

cl_vector_sizes = {2, 3, 4, 8, 16}

np_c_dtypes = {dtype(n): c for n, c in zip(np_dtypes, c_dtypes)}


def dtype_decl(dt):
    """OpenCL-C spelling of a numpy dtype ('float', 'double3', ...); kept because
    it is part of the reference's helper surface (tests/test_misc.py upstream)."""
    dt = dtype(dt)
    if not dt.shape:
        return np_c_dtypes[dt]
    if len(dt.shape) != 1:
        raise ValueError("Too many vector dimensions: {}".format(dt.shape))
    n, = dt.shape
    if n not in cl_vector_sizes:
        raise ValueError("Invalid vector size: {}".format(n))
    return "{}{}".format(np_c_dtypes[dt.base], n)


def dtype_sizeof(dt):
    """Bytes one element occupies on the device: a 3-vector fills a 4-lane slot."""
    dt = dtype(dt)
    if dt.base in np_c_dtypes:
        *outer, n = dt.shape or (1,)
        if n != 1 and n not in cl_vector_sizes:
            raise ValueError("Invalid vector size: {}".format(n))
        return dt.base.itemsize * product(outer) * (4 if n == 3 else n)
    subtype, shape = dt.subdtype
    return product(shape) * dtype_sizeof(subtype)


def dtype_lanes(dt):
    """Scalars per device slot of a scalar / vector dtype (3 -> 4)."""
    dt = dtype(dt)
    return dtype_sizeof(dt) // dt.base.itemsize
