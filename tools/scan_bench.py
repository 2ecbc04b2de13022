"""Standalone exclusive-scan timing (BASELINE.json configs[2]): python tools/scan_bench.py [--n 67108864]"""
import argparse, sys
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from collision_b200 import _lib
ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1 << 26)
ap.add_argument("--reps", type=int, default=7)
args = ap.parse_args()
dev = torch.device("cuda", 0)
L, P = _lib.lib(), _lib.ptr
S = torch.cuda.current_stream().cuda_stream
n = args.n
g = torch.Generator(device=dev).manual_seed(4)
vals = torch.randint(0, 128, (n,), dtype=torch.int32, device=dev, generator=g)
ref = torch.cumsum(vals.to(torch.int64), 0) - vals
work = vals.clone()
scr = _lib.alloc_bytes(dev, L.col_scan_scratch_bytes(n))
tot = torch.zeros(1, dtype=torch.int64, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ts = []
for _ in range(args.reps + 2):
    work.copy_(vals); flush.fill_(1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); rc = L.col_exclusive_scan_u32(P(work), n, P(tot), P(scr), S); b.record(); b.synchronize()
    assert rc == 0
    ts.append(a.elapsed_time(b))
ms = float(np.median(ts[2:]))
ok = torch.equal(work.to(torch.int64) & 0xFFFFFFFF, ref & 0xFFFFFFFF) and int(tot.item()) == int(ref[-1] + vals[-1])
print("scan n={} {:.3f} ms {:.1f} GB/s ({})".format(n, ms, 8 * n / ms / 1e6, "ok" if ok else "MISMATCH"))
