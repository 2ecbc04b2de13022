"""numpy model of the host-checkable logic of the sharded build (mode "sharded_build"):

  * the PLAN of the sort by exchange (csrc/sort.cu: plan_kernel, collect_push_kernel,
    col_shard_sort_received): top-digit counts per slice -> bucket starts -> for every rank the buckets
    covering its positions, where each source's entries start in its receive buffers -> what a rank
    receives (input order) -> its stable sort, placed at `offset`: must equal the same positions of the
    stable sort of the whole scene;
  * the RANGE TEST of the two-level top pass (csrc/refit.cu: top_local_kernel): for a node [lo, hi]
    inside a rank's range of leaves [rb, re), "the parent lies inside the range too" is decided from the
    node alone -- delta(hi, re) < delta(hi, hi + 1) for a left child, delta(rb - 1, lo) < delta(lo - 1, lo)
    for a right child -- checked against the radix tree built top-down from the sorted keys; the nodes where
    it fails are the rank's range roots: they tile the range and are few.

Same arithmetic as the kernels (delta as collision.cl:55-70 for in-range arguments, ties broken by
position).  Development aid and CPU test (tests/test_shard_model.py); not imported by the package."""
import numpy as np

RADIX = 256


# ------------------------------------------------------------------ the plan of the exchange
def plan(hist_all, pos, n, me):
    """hist_all: world x 256 counts; pos: world x (pos_lo, pos_hi), pos_lo > pos_hi = nothing.
    Returns world x [offset, count, first bucket, last bucket, base of `me`'s entries]."""
    world = hist_all.shape[0]
    total = hist_all.sum(axis=0)
    start = np.concatenate([[0], np.cumsum(total)])          # start[256] = n
    before_me = hist_all[:me].sum(axis=0)
    out = np.zeros((world, 5), np.int64)
    for r in range(world):
        lo, hi = pos[r]
        hi = min(hi, n - 1)
        if lo > hi or lo >= n:
            out[r] = [0, 0, RADIX, RADIX, 0]
            continue
        # the bucket that holds position p = the last one that starts at or before p
        d_lo = max(d for d in range(RADIX) if start[d] <= lo)
        d_hi = max(d for d in range(RADIX) if start[d] <= hi)
        out[r] = [start[d_lo], start[d_hi + 1] - start[d_lo], d_lo, d_hi, before_me[d_lo:d_hi + 1].sum()]
    return out


def exchange(codes, world, end_bit=30):
    """Every rank's part of the sorted order as the exchange produces it.  Returns per rank
    (offset, sorted codes, sorted ids) and checks the receive-buffer bookkeeping on the way."""
    n = len(codes)
    shift = end_bit - 8 if end_bit > 8 else 0
    mask = 0xFF if end_bit > 8 else (1 << end_bit) - 1
    cuts = [n * r // world for r in range(world + 1)]
    digit = (codes >> shift) & mask
    hist_all = np.stack([np.bincount(digit[cuts[s]:cuts[s + 1]], minlength=RADIX) for s in range(world)])
    pos = [(max(cuts[r], 1) - 1, min(cuts[r + 1], n - 1)) if cuts[r + 1] > cuts[r] else (1, 0) for r in range(world)]
    plans = [plan(hist_all, pos, n, me) for me in range(world)]
    out = []
    for r in range(world):
        off, cnt, d_lo, d_hi = plans[0][r][:4]
        assert all((plans[s][r][:4] == plans[0][r][:4]).all() for s in range(world))     # the same plan everywhere
        recv_codes = np.full(cnt, -1, np.int64)
        recv_ids = np.full(cnt, -1, np.int64)
        for s in range(world):      # every source compacts its own slice, in order, to its base
            own = np.arange(cuts[s], cuts[s + 1])
            sel = own[(digit[own] >= d_lo) & (digit[own] <= d_hi)] if cnt else own[:0]
            base = plans[s][r][4]
            assert (recv_ids[base:base + len(sel)] == -1).all()                          # nobody else writes there
            recv_codes[base:base + len(sel)] = codes[sel]
            recv_ids[base:base + len(sel)] = sel
        assert (recv_ids >= 0).all() and (np.diff(recv_ids) > 0).all()                    # complete, in input order
        order = np.argsort(recv_codes & ((1 << end_bit) - 1), kind="stable")
        if cnt:
            assert off <= pos[r][0] and pos[r][1] < off + cnt
        out.append((int(off), recv_codes[order], recv_ids[order]))
    return out


# ------------------------------------------------------------------ the range test of the local top pass
def clz32(x):
    return 32 - int(x).bit_length()


def key_delta(a, code_a, b, code_b):
    x = int(code_a) ^ int(code_b)
    return clz32(x) if x else 32 + clz32(a ^ b)


def radix_tree(codes):
    """Internal nodes of the radix tree over the sorted (code, position) keys, top-down: a node
    [lo, hi] splits where the prefix shared by neighbours is shortest.  Returns {(lo, hi): parent (lo, hi)}."""
    n = len(codes)
    d = [key_delta(i, codes[i], i + 1, codes[i + 1]) for i in range(n - 1)]
    parent = {}
    stack = [(0, n - 1)]
    while stack:
        lo, hi = stack.pop()
        if lo == hi:
            continue
        g = lo + int(np.argmin(d[lo:hi]))
        for child in ((lo, g), (g + 1, hi)):
            parent[child] = (lo, hi)
            stack.append(child)
    return parent, d


def parent_inside(lo, hi, codes, d, rb, re):
    """top_local_kernel's test, for a node [lo, hi] inside [rb, re)."""
    n = len(codes)
    dl = d[lo - 1] if lo > 0 else -1
    dr = d[hi] if hi < n - 1 else -1
    if dr > dl:      # a left child: the parent extends to the right
        return re >= n or key_delta(hi, codes[hi], re, codes[re]) < dr
    return rb == 0 or key_delta(rb - 1, codes[rb - 1], lo, codes[lo]) < dl


def check_ranges(codes, world):
    """For every rank's range: the test agrees with the tree for every node inside the range, and the
    nodes where it fails (range roots) tile the range.  Returns the number of range roots per rank."""
    n = len(codes)
    parent, d = radix_tree(codes)
    cuts = [n * r // world for r in range(world + 1)]
    counts = []
    for r in range(world):
        rb, re = cuts[r], cuts[r + 1]
        roots = []
        for (lo, hi), (plo, phi) in parent.items():     # (leaves included: lo == hi)
            if lo < rb or hi >= re:
                continue
            truth = rb <= plo and phi < re
            assert parent_inside(lo, hi, codes, d, rb, re) == truth, (lo, hi, plo, phi, rb, re)
            if not truth:
                roots.append((lo, hi))
        roots.sort()
        covered = [p for lo, hi in roots for p in range(lo, hi + 1)]
        if re > rb and not (rb == 0 and re == n):
            assert covered == list(range(rb, re)), (r, roots[:5])
        counts.append(len(roots))
    return counts


def self_test():
    rs = np.random.RandomState(7)
    for n, world, kind in [(1, 1, "uniform"), (5, 3, "uniform"), (2000, 4, "uniform"), (3000, 8, "few"),
                           (2500, 5, "one_bucket"), (4096, 2, "clustered"), (1500, 7, "equal")]:
        if kind == "uniform":
            codes = rs.randint(0, 1 << 30, n)
        elif kind == "few":
            codes = rs.randint(0, 7, n) * 0x01234567 % (1 << 30)
        elif kind == "one_bucket":
            codes = (5 << 22) | rs.randint(0, 1 << 22, n)
        elif kind == "clustered":
            codes = rs.randint(0, 3, n) * (300 << 20) + rs.randint(0, 1 << 21, n)
        else:
            codes = np.full(n, 12345)
        codes = codes.astype(np.int64)
        order = np.argsort(codes, kind="stable")
        for r, (off, sc, si) in enumerate(exchange(codes, world)):
            assert (sc == codes[order][off:off + len(sc)]).all(), (kind, n, world, r)
            assert (si == order[off:off + len(si)]).all(), (kind, n, world, r)
        if n >= 2:
            counts = check_ranges(codes[order], world)
            # at most two range roots per level of the tree (30 code bits + the tie-break bits)
            assert max(counts) <= 2 * (30 + int(n).bit_length()), counts
    return True


if __name__ == "__main__":
    print("ok" if self_test() else "failed")
