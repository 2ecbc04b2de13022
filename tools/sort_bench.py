"""Standalone radix-sort / scan timing (BASELINE.json configs[2]) -- development aid and ncu target.
    python tools/sort_bench.py [--n 67108864] [--reps 5] [--pairs]
"""
import argparse
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from collision_b200 import _lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1 << 26)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--pairs", action="store_true")
    ap.add_argument("--dist", default="uniform32")
    ap.add_argument("--check", action="store_true")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    L, P = _lib.lib(), _lib.ptr
    S = torch.cuda.current_stream().cuda_stream
    n = args.n
    g = torch.Generator(device=dev).manual_seed(4)
    if args.dist == "uniform32":
        keys = torch.randint(-2**31, 2**31 - 1, (n,), dtype=torch.int32, device=dev, generator=g)
    elif args.dist == "uniform30":
        keys = torch.randint(0, 2**30, (n,), dtype=torch.int32, device=dev, generator=g)
    else:
        keys = torch.arange(n, dtype=torch.int32, device=dev)
    vb = 4 if args.pairs else 0
    scratch = _lib.alloc_bytes(dev, L.col_sort_scratch_bytes(n, 4, vb))
    kout = torch.empty_like(keys)
    vin = torch.arange(n, dtype=torch.int32, device=dev) if args.pairs else None
    vout = torch.empty_like(keys) if args.pairs else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for _ in range(args.reps + 2):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        rc = L.col_sort_pairs(P(keys), P(kout), P(vin), P(vout), n, 4, vb, 0, P(scratch), S)
        b.record()
        b.synchronize()
        assert rc == 0, rc
        ts.append(a.elapsed_time(b))
    ms = float(np.median(ts[2:]))
    nbytes = (68 if args.pairs else 36) * n
    print("sort {} n={} {}: {:.3f} ms  {:.1f} GB/s algorithmic  {:.2f} Gkeys/s".format(
        "pairs" if args.pairs else "keys", n, args.dist, ms, nbytes / ms / 1e6, n / ms / 1e6))
    if args.check:
        ref = torch.sort(keys.to(torch.int64) & 0xFFFFFFFF, stable=True)
        got = kout.to(torch.int64) & 0xFFFFFFFF
        assert torch.equal(got, ref.values)
        if args.pairs:
            assert torch.equal(vout.to(torch.int64), ref.indices)
        print("check ok")


if __name__ == "__main__":
    main()
