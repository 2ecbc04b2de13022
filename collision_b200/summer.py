"""Vector sum (mirror of collision/summer.py upstream).  Floating-point sums
depend on the association order, so they are compared with a tolerance
(upstream's own test uses assert_almost_equal), not bit for bit; integer sums
(upstream tests uint32x4 and int32, tests/test_summer_py.py:12) wrap and are
exact in any order."""
from .reduce import ReductionProgram, Reducer


class SumProgram(ReductionProgram):
    accumulator = [("0", "ADD")]                                # collision/summer.py:5
    _entry = "col_sum"
    _value_bases = ('float32', 'float64', 'uint32', 'int32', 'uint64', 'int64')


class Summer(Reducer):
    program_type = SumProgram
