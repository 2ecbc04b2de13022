"""Development aid (torchrun, 2+ GPUs): one ShardedCollider(mode=...) call on a small scene, rows checked
against the single-GPU Collider on every rank.  python -m torch.distributed.run ... tools/shard_debug.py [--spheres N] [--mode M]"""
import argparse
import datetime
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spheres", type=int, default=20011)
    ap.add_argument("--mode", default="sharded_build")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--sort", default=None, choices=["exchange", "local", "replicated"])
    ap.add_argument("--one-device", action="store_true", help="every rank on cuda:0, gloo for the collectives (cheap multi-rank "
                    "check of the kernels and the peer mappings on a one-GPU box; no graph capture)")
    ap.add_argument("--tiny-send", action="store_true", help="start with a root exchange that is far too small (the call must repeat itself)")
    ap.add_argument("--graph", action="store_true", help="sync_totals=False calls: plain, plain, capture, replay ...")
    args = ap.parse_args()
    import faulthandler
    faulthandler.dump_traceback_later(90, exit=True)      # a hang shows where
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", 0 if args.one_device else int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    if args.one_device:
        dist.init_process_group("gloo", timeout=datetime.timedelta(seconds=60))
    else:
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=60))
    from collision_b200.collision import Collider
    from collision_b200.sharded import ShardedCollider
    n = args.spheres
    g = torch.Generator(device=dev).manual_seed(5)
    coords = torch.zeros(n, 4, device=dev)
    coords[:, :3] = torch.rand(n, 3, device=dev, generator=g)
    radii = torch.full((n,), 0.5 * n ** (-1.0 / 3.0), device=dev)
    cq = torch.cuda.Stream(dev)
    one = Collider(dev, n, 8, 128)
    nb = torch.zeros(1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    one.get_collisions(cq, coords, radii, nb, None, 0).synchronize()
    total = int(nb.item()) & 0xFFFFFFFF
    want = torch.empty(max(total, 1) * 2, dtype=torch.int32, device=dev)
    one.get_collisions(cq, coords, radii, nb, want, total).synchronize()
    print("rank {} single-GPU total {}".format(rank, total), flush=True)
    col = ShardedCollider(dev, n, 8, 128, mode=args.mode, graph=args.graph, sharded_sort=args.sort)
    if args.tiny_send:
        col._send_cap = 16
    print("rank {} sharded collider up: queries [{}, {})".format(rank, col.q_begin, col.q_end), flush=True)
    for again in range(2):
        cc0, rr0 = coords.clone(), radii.clone()
        torch.cuda.synchronize()        # (the clones are made on the default stream)
        col.get_collisions(cq, cc0, rr0, nb, None, 0).synchronize()
        print("rank {} count-only: local {} base {} total {} (totals {}, root exchange {})".format(
            rank, col.local_pairs, col.pair_base, col.total_pairs, col._totals.tolist(), getattr(col, "_send_cap", None)), flush=True)
    cap = col.local_pairs
    rows = torch.empty(max(cap, 1) * 2, dtype=torch.int32, device=dev)
    cc, rr = coords.clone(), radii.clone()
    if args.mode in ("gather_inputs", "sharded_build"):     # the rank holds its own slice only: poison the rest
        from collision_b200.sharded import shard_range
        b, e = shard_range(n, world, rank)
        coords, radii = coords.clone(), radii.clone()
        coords[:b] = float("nan"); coords[e:] = float("nan"); radii[:b] = float("nan"); radii[e:] = float("nan")
    for step in range(args.steps):
        if args.graph:
            cc.copy_(coords); rr.copy_(radii); torch.cuda.synchronize()
            print("rank {} step {} enqueue (graphs {}, replays {})".format(rank, step, len(col._graphs), col.graph_replays), flush=True)
            col.get_collisions(cq, cc, rr, nb, rows, cap, sync_totals=False).synchronize()
            t = col._totals.tolist()
            col.total_pairs, col.pair_base = sum(t), sum(t[:rank])
        else:
            cc1, rr1 = coords.clone(), radii.clone()
            torch.cuda.synchronize()
            col.get_collisions(cq, cc1, rr1, nb, rows, cap).synchronize()
        t = col._totals.tolist()
        mine = min(t[rank], cap)
        got = rows[:mine * 2].view(-1, 2)
        ref = want[col.pair_base * 2:(col.pair_base + mine) * 2].view(-1, 2)
        ok = col.total_pairs == total and t[rank] == cap and got.shape == ref.shape and bool((got == ref).all())
        print("rank {} step {} rows equal: {} (totals {}, count-only said {})".format(rank, step, ok, t, cap), flush=True)
        if not ok and got.shape == ref.shape:
            bad = (got != ref).any(dim=1).nonzero().flatten()
            print("rank {}: {} rows differ, first at {}: got {} want {}; next rows got {} want {}".format(
                rank, bad.numel(), int(bad[0]) if bad.numel() else -1,
                got[bad[0]].tolist() if bad.numel() else None, ref[bad[0]].tolist() if bad.numel() else None,
                got[bad[0]:bad[0] + 3].tolist() if bad.numel() else None,
                ref[bad[0]:bad[0] + 3].tolist() if bad.numel() else None), flush=True)
    if args.mode == "sharded_build":
        lst = col._list.view(torch.int32)
        print("rank {} root exchange cut to {} of {}".format(rank, col._send_cap, col._cap), flush=True)
        xh = col._xchg.view(torch.int32)[:4].tolist()
        print("rank {} range roots {}, spanning nodes listed {} (flags {} {})".format(rank, xh[0], xh[2], xh[1], xh[3]), flush=True)
        print("rank {} roots packed {} of cap {}, remote subtrees fetched {} of {}".format(
            rank, int(col._packed.view(torch.int32)[0].item()), col._cap, int(lst[0].item()), col._list_cap), flush=True)
    col.close()
    print("rank {} closed".format(rank), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    print("rank {} process group destroyed".format(rank), flush=True)


if __name__ == "__main__":
    main()
