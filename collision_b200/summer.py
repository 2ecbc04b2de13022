"""Vector sum (mirror of collision/summer.py upstream).  Floating-point sums
depend on the association order, so this one is compared with a tolerance
(upstream's own test uses assert_almost_equal), not bit for bit."""
from .reduce import ReductionProgram, Reducer


class SumProgram(ReductionProgram):
    accumulator = [("0", "ADD")]                                # collision/summer.py:5
    _entry = "col_sum"


class Summer(Reducer):
    program_type = SumProgram
