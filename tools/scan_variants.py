"""A/B timing of the exclusive scan in differently built libcollision_b200.so files (development aid):
    python tools/scan_variants.py lib_a.so lib_b.so ... [--n 67108864 16777216] [--reps 9]"""
import argparse
import ctypes as C
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from collision_b200 import _lib  # noqa: E402


def load(path):
    h = C.CDLL(str(path))
    for name in ("col_exclusive_scan_u32", "col_scan_scratch_bytes"):
        res, args = _lib.SIGNATURES[name]
        fn = getattr(h, name)
        fn.restype, fn.argtypes = res, args
    return h


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("libs", nargs="+")
    ap.add_argument("--n", type=int, nargs="+", default=[1 << 26, 1 << 24])
    ap.add_argument("--reps", type=int, default=9)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    P = _lib.ptr
    S = torch.cuda.current_stream().cuda_stream
    libs = [(Path(p).stem, load(p)) for p in args.libs]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for n in args.n:
        g = torch.Generator(device=dev).manual_seed(4)
        vals = torch.randint(0, 128, (n,), dtype=torch.int32, device=dev, generator=g)
        ref = torch.cumsum(vals.to(torch.int64), 0) - vals
        work = vals.clone()
        tot = torch.zeros(1, dtype=torch.int64, device=dev)
        times = {name: [] for name, _ in libs}
        for rep in range(args.reps + 2):
            for name, L in libs:
                scr = _lib.alloc_bytes(dev, L.col_scan_scratch_bytes(n))
                work.copy_(vals)
                flush.fill_(1)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                rc = L.col_exclusive_scan_u32(P(work), n, P(tot), P(scr), S)
                b.record()
                b.synchronize()
                assert rc == 0, (name, rc)
                if rep == 0:
                    ok = torch.equal(work.to(torch.int64) & 0xFFFFFFFF, ref & 0xFFFFFFFF) and \
                        int(tot.item()) == int(ref[-1] + vals[-1])
                    assert ok, name
                times[name].append(a.elapsed_time(b))
        for name, _ in libs:
            t = np.array(times[name][2:])
            print("n={:10d} {:12s} median {:8.2f} us  min {:8.2f} us  {:6.1f} GB/s (8 B/value)".format(
                n, name, 1e3 * float(np.median(t)), 1e3 * float(t.min()), 8 * n / float(np.median(t)) / 1e6), flush=True)


if __name__ == "__main__":
    main()
