"""Probe (2+ GPUs, torchrun): collision_b200.shared.SharedBuffer -- can this rank's kernels read a
peer's buffer in place, and how fast?"""
import datetime
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
    from collision_b200 import _lib
    from collision_b200.shared import SharedBuffer
    n = 256 << 20
    sb = SharedBuffer(dev, n)
    sb.local.fill_(rank + 1)
    torch.cuda.synchronize()
    dist.barrier()
    nxt = (rank + 1) % world
    view = sb.peer(nxt)
    print("rank {} reads peer {}: first byte {} (want {})".format(rank, nxt, int(view[123].item()), nxt + 1), flush=True)
    dst = torch.empty(n, dtype=torch.uint8, device=dev)
    L = _lib.lib()
    s = torch.cuda.current_stream()
    for _ in range(2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = L.col_copy(dst.data_ptr(), sb.peer_ptrs[nxt], n, s.cuda_stream)
        e1.record(); e1.synchronize()
    ok = rc == 0 and bool((dst == nxt + 1).all())
    print("rank {} col_copy from the peer pointer: ok={} {:.1f} GB/s".format(rank, ok, n / e0.elapsed_time(e1) / 1e6), flush=True)
    dist.barrier()
    sb.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
