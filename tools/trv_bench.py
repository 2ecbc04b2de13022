"""Traversal-only timing on a built scene (development aid): python tools/trv_bench.py [--n N] [--clustered]"""
import argparse, sys
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from collision_b200 import _lib
from collision_b200.collision import Collider
sys.path.insert(0, str(ROOT / "tools"))
from stage_bench import gen_uniform, gen_clustered, Timer

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1 << 20)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--clustered", action="store_true")
ap.add_argument("--no-flush", action="store_true")
ap.add_argument("--two-launch-build", action="store_true")
args = ap.parse_args()
dev = torch.device("cuda", 0)
L, P = _lib.lib(), _lib.ptr
cq = torch.cuda.current_stream(); S = cq.cuda_stream
n = args.n
coords, radii = (gen_clustered if args.clustered else gen_uniform)(n, dev)
col = Collider(dev, n, 8, 128, fused_build=not args.two_launch_build)
n_buf = torch.zeros(1, dtype=torch.int32, device=dev)
col.get_collisions(cq, coords, radii, n_buf, None, 0).synchronize()
m = int(n_buf.item()) & 0xFFFFFFFF
pairs = torch.empty(max(m, 1) * 2, dtype=torch.int32, device=dev)
scr_ = _lib.alloc_bytes(dev, L.col_traverse_scratch_bytes(n, m))
T = Timer(dev, args.reps)
if args.no_flush:
    T.flush = torch.empty(16, dtype=torch.uint8, device=dev)
code = 0
t1 = T(lambda: L.col_traverse_wide_count(P(col._total_buf), None, P(col._wide_buf), P(col._bounds_buf), P(col._scene_buf), code, n, 0, n, S))
t2 = T(lambda: L.col_traverse_wide_pairs(P(pairs), m, 0, P(col._total_buf), P(col._wide_buf), P(col._bounds_buf), P(col._scene_buf), P(col._ids_bufs[1]), code, n, 0, n, P(scr_), S))
t3 = T(lambda: col.get_collisions(cq, coords, radii, n_buf, pairs, m))
print("n={} m={} count {:.3f} ms  pairs {:.3f} ms  full {:.3f} ms".format(n, m, t1, t2, t3))
