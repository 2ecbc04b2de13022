"""Times col_shard_sort (one rank's part of the sort of ALL codes) beside the full col_sort_pairs, one GPU.
python tools/shard_sort_bench.py [--log2 27] [--world 8]"""
import argparse
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in evs:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in evs)
    return t[len(t) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2", type=int, default=27)
    ap.add_argument("--world", type=int, default=8)
    args = ap.parse_args()
    from collision_b200 import _lib
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    L, P = _lib.lib(), _lib.ptr
    n = 1 << args.log2
    S = torch.cuda.current_stream(dev).cuda_stream
    g = torch.Generator(device=dev).manual_seed(1)
    codes = torch.randint(0, 1 << 30, (n,), dtype=torch.int32, device=dev, generator=g)
    ids = torch.arange(n, dtype=torch.int32, device=dev)
    oc, oi = torch.empty_like(codes), torch.empty_like(ids)
    full = _lib.alloc_bytes(dev, L.col_sort_scratch_bytes(n, 4, 4))
    part = _lib.alloc_bytes(dev, L.col_shard_sort_scratch_bytes(n))
    rng = torch.zeros(4, dtype=torch.int32, device=dev)
    t_full = timed(lambda: L.col_sort_pairs(P(codes), P(oc), P(ids), P(oi), n, 4, 4, 30, P(full), S))
    want_c, want_i = oc.clone(), oi.clone()
    print("full sort of 2^{} pairs: {:.3f} ms".format(args.log2, t_full))
    for r in (0, args.world // 2, args.world - 1):
        lo, hi = max(n * r // args.world, 1) - 1, min(n * (r + 1) // args.world, n - 1)
        oc.fill_(-1)
        oi.fill_(-1)
        t = timed(lambda: L.col_shard_sort(P(codes), n, lo, hi, 30, P(oc), P(oi), P(rng), P(part), S))
        off, cnt = rng.tolist()[:2]
        ok = torch.equal(oc[off:off + cnt], want_c[off:off + cnt]) and torch.equal(oi[off:off + cnt], want_i[off:off + cnt])
        print("rank {} of {}: positions [{}, {}] -> {} pairs collected, {:.3f} ms ({:.2f}x), equal to the full sort there: {}".format(
            r, args.world, lo, hi, cnt, t, t_full / t, ok))


if __name__ == "__main__":
    main()
