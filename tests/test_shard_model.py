"""Host-side check of the logic behind mode "sharded_build" on its numpy model (tools/shard_model.py): the plan of
the sort by exchange (bucket selection, receive-buffer bases, input order => bit-identical positions) and the
range test of the two-level top pass against a radix tree built top-down (range roots tile a rank's range and
are few; duplicates and single-bucket scenes included).  No GPU needed."""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tools"))


def test_model_self_test():
    import shard_model
    assert shard_model.self_test()


def test_plan_of_an_empty_rank_and_of_one_bucket():
    import shard_model as M
    hist = np.zeros((3, 256), np.int64)
    hist[0, 9] = 4
    hist[1, 9] = 5
    hist[2, 9] = 1
    # everything in bucket 9: every rank with positions takes all of it, in rank order of the sources
    pos = [(0, 4), (1, 0), (4, 9)]
    for me, base in ((0, 0), (1, 4), (2, 9)):
        p = M.plan(hist, pos, 10, me)
        assert p[0].tolist() == [0, 10, 9, 9, base]
        assert p[1].tolist() == [0, 0, 256, 256, 0]          # the rank without positions takes nothing
        assert p[2].tolist() == [0, 10, 9, 9, base]


def test_range_roots_are_few_on_a_larger_scene():
    import shard_model as M
    rs = np.random.RandomState(11)
    codes = np.sort(rs.randint(0, 1 << 30, 20000).astype(np.int64), kind="stable")
    counts = M.check_ranges(codes, 8)
    assert max(counts) <= 64 and min(counts) >= 1, counts
