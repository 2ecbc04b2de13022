"""A/B timing of differently built libcollision_b200.so files on the same keys (development aid):
    python tools/sort_variants.py lib_a.so lib_b.so ... [--n 1048576 2097152] [--reps 15]
Builds come from `COL_NVCC_EXTRA="-D..." python -m collision_b200.build --force` (see build.py)."""
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
    for name in ("col_sort_pairs", "col_sort_scratch_bytes"):
        res, args = _lib.SIGNATURES[name]
        fn = getattr(h, name)
        fn.restype, fn.argtypes = res, args
    return h


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("libs", nargs="+")
    ap.add_argument("--n", type=int, nargs="+", default=[1 << 20, 1 << 21])
    ap.add_argument("--reps", type=int, default=15)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    P = _lib.ptr
    S = torch.cuda.current_stream().cuda_stream
    libs = [(Path(p).stem, load(p)) for p in args.libs]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for n in args.n:
        g = torch.Generator(device=dev).manual_seed(4)
        keys = torch.randint(0, 2**30, (n,), dtype=torch.int32, device=dev, generator=g)
        vin = torch.arange(n, dtype=torch.int32, device=dev)
        ref = torch.sort(keys, stable=True)
        for end_bit in (30, 32):
            times = {name: [] for name, _ in libs}
            for rep in range(args.reps + 2):
                for name, L in libs:
                    scratch = _lib.alloc_bytes(dev, L.col_sort_scratch_bytes(n, 4, 4))
                    kout, vout = torch.empty_like(keys), torch.empty_like(keys)
                    flush.fill_(1)
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    rc = L.col_sort_pairs(P(keys), P(kout), P(vin), P(vout), n, 4, 4, end_bit, P(scratch), S)
                    b.record()
                    b.synchronize()
                    assert rc == 0, rc
                    if rep == 0:
                        assert torch.equal(kout, ref.values) and torch.equal(vout.long(), ref.indices), name
                    times[name].append(a.elapsed_time(b))
            for name, _ in libs:
                t = np.array(times[name][2:])
                print("n={:9d} bits={} {:10s} median {:7.2f} us  min {:7.2f} us".format(
                    n, end_bit, name, 1e3 * float(np.median(t)), 1e3 * float(t.min())), flush=True)


if __name__ == "__main__":
    main()
