"""ShardedCollider(balance=True) on a skewed scene, frame after frame (run under torchrun, >= 2 GPUs):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/balance_bench.py [--spheres 4194304] [--frames 6]
Half the spheres sit in a few dense clusters, half are spread out, so equal query counts per rank mean
unequal rows.  Every frame prints the cut, the rows per rank and the step time (max over ranks, CUDA
events); the rows of all ranks together are checked against the first frame's count."""
import argparse
import datetime
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from collision_b200.sharded import ShardedCollider  # noqa: E402


def skewed_scene(n, dev, seed=4):
    g = torch.Generator(device=dev).manual_seed(seed)
    c = torch.zeros(n, 4, device=dev)
    half = n // 2
    centres = torch.rand(64, 3, device=dev, generator=g) * 0.5              # dense clusters in one octant
    which = torch.randint(0, 64, (half,), device=dev, generator=g)
    c[:half, :3] = centres[which] + torch.randn(half, 3, device=dev, generator=g) * 0.03
    c[half:, :3] = torch.rand(n - half, 3, device=dev, generator=g)
    r = torch.full((n,), 0.5 * n ** (-1.0 / 3.0), device=dev)
    return c, r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spheres", dest="n", type=int, default=1 << 22)
    ap.add_argument("--frames", type=int, default=6)
    ap.add_argument("--mode", default="replicate_build")
    ap.add_argument("--scene", choices=["skewed", "configs3"], default="skewed",
                    help="configs3: BASELINE.json configs[3], 1024 Gaussian clusters (use --spheres 16777216)")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
    n = args.n
    if args.scene == "configs3":                  # the host generator of bench.py / the full-size test: identical on every rank
        sys.path.insert(0, str(ROOT))
        from bench import clustered_scene
        c, r = clustered_scene(n)
        coords, radii = torch.from_numpy(c).to(dev), torch.from_numpy(r).to(dev)
    else:
        coords, radii = skewed_scene(n, dev)      # same seed on every rank: identical scene
    cq = torch.cuda.Stream(dev)
    n_buf = torch.zeros(1, dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for balance in (False, True):
        col = ShardedCollider(dev, n, 8, 128, mode=args.mode, balance=balance)
        torch.cuda.synchronize(dev)
        cq.wait_stream(torch.cuda.current_stream(dev))
        col.get_collisions(cq, coords, radii, n_buf, None, 0).synchronize()      # count only: sizes the buffers
        total = col.total_pairs
        # any cut must fit: room for the larger of an equal share x 2 and what the first cut needs
        cap = max(2 * total // world, max(col._totals.tolist())) + 1024
        pairs = torch.empty(cap * 2, dtype=torch.int32, device=dev)
        for frame in range(args.frames):
            edges = list(col.edges)
            flush.fill_(1)
            dist.barrier()
            torch.cuda.synchronize(dev)
            cq.wait_stream(torch.cuda.current_stream(dev))
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(cq)
            col.get_collisions(cq, coords, radii, n_buf, pairs, cap, sync_totals=False)
            b.record(cq)
            b.synchronize()
            ms = torch.tensor([a.elapsed_time(b)], device=dev)
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            col.get_collisions(cq, coords, radii, n_buf, pairs, cap).synchronize()   # untimed: host bookkeeping + re-cut
            rows = col._totals.tolist()
            assert sum(rows) == total and max(rows) <= cap, (rows, total, cap)
            if rank == 0:
                print("balance={} frame {} step {:.3f} ms  rows/rank {}  max/mean {:.2f}  cut {}".format(
                    balance, frame, float(ms.item()), rows, max(rows) * world / max(total, 1),
                    edges if world <= 4 else "..."), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
