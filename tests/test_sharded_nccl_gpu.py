"""ShardedCollider.get_collisions on REAL GPUs over NCCL (2 ranks, one process per GPU): every
mode, balance=True over several frames and the CUDA-graph replay of the whole step, each rank's
rows concatenated by pair_base and compared bit for bit with the oracle's single-rank list
(= the reference's work-items run sequentially).  Skipped on boxes with fewer than 2 GPUs
(`gpurun --gpus 2 -- python -m pytest tests/test_sharded_nccl_gpu.py -m gpu`)."""
import os
import socket

import numpy as np
import pytest

import golden_data as G

pytestmark = pytest.mark.gpu

# "sharded_build" alone = the default: the sort dealt out as well, spheres pushed to the ranks that need them;
# "/local_sort" = codes and scene all-gathered, every rank sorts its own buckets; "/replicated_sort" = ... sorts everything
MODES = ("broadcast_tree", "replicate_build", "gather_inputs", "sharded_build", "sharded_build/local_sort",
         "sharded_build/replicated_sort")
SORT_VARIANTS = {"": "exchange", "local_sort": "local", "replicated_sort": "replicated", "tiny_send": "exchange"}


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _scene(n, kind):
    c, r = G.random_spheres(n)
    c = np.ascontiguousarray(c, dtype=np.float32).copy()
    if kind == "skewed":          # half the spheres packed into a corner: unequal rows per rank
        c[: n // 2, :3] *= 0.2
    if kind == "duplicates":      # runs of equal Morton codes straddling the rank boundary
        c[n // 3: n // 3 + 700, :3] = c[n // 3, :3]
    return c, np.ascontiguousarray(r, dtype=np.float32)


def _worker(rank, world, port, n, kind, mode, balance, graph, frames, raw, tmpdir, one_device=False):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = torch.device("cuda", 0 if one_device else rank)
    torch.cuda.set_device(dev)
    if one_device:      # every rank on the same GPU: gloo moves the collectives' bytes, CUDA IPC maps the peers' buffers
        dist.init_process_group("gloo", rank=rank, world_size=world)
    else:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from collision_b200.sharded import ShardedCollider, shard_range
        mode, _, variant = mode.partition("/")
        c, r = _scene(n, kind)
        b, e = shard_range(n, world, rank)
        cq = torch.cuda.Stream(dev)

        def inputs():
            """What the rank holds before a call: everything on rank 0 for the broadcast modes,
            its own slice of the spheres for the sharded-input modes; the rest is poison."""
            if mode in ("gather_inputs", "sharded_build"):
                cc, rr = np.full_like(c, np.nan), np.full_like(r, np.nan)
                cc[b:e], rr[b:e] = c[b:e], r[b:e]
            elif rank == 0:
                cc, rr = c, r
            else:
                cc, rr = np.full_like(c, np.nan), np.full_like(r, np.nan)
            ct, rt = torch.from_numpy(cc.copy()).to(dev), torch.from_numpy(rr.copy()).to(dev)
            if raw:
                ct, rt = ct.view(torch.uint8).reshape(-1), rt.view(torch.uint8).reshape(-1)
            return ct, rt

        col = ShardedCollider(dev, n, 8, 64, mode=mode, balance=balance, graph=graph,
                              sharded_sort=SORT_VARIANTS[variant])
        assert col.sort_mode == (SORT_VARIANTS[variant] if mode == "sharded_build" else "replicated")
        if variant == "tiny_send":      # a root exchange cut down too far: the call notices, resets it and repeats itself
            col._send_cap = 16
        n_buf = torch.zeros(1, dtype=torch.int32, device=dev)
        ct, rt = inputs()
        torch.cuda.synchronize()
        col.get_collisions(cq, ct, rt, n_buf, None, 0).synchronize()          # the sizing call
        cap = col.suggested_capacity(3.0) if balance else col.local_pairs
        assert col.local_pairs <= cap
        rows = torch.full((max(cap, 1) * 2,), -1, dtype=torch.int32, device=dev)
        log = []
        for f in range(frames):
            ct, rt = inputs() if not graph else (ct, rt)
            if graph and f:      # same buffers, refilled: the capture must read the new contents
                fresh_c, fresh_r = inputs()
                ct.copy_(fresh_c)
                rt.copy_(fresh_r)
            rows.fill_(-1)
            torch.cuda.synchronize()
            if graph:
                col.get_collisions(cq, ct, rt, n_buf, rows, cap, sync_totals=False).synchronize()
                totals = col._totals.tolist()
                local, base, total = totals[rank], sum(totals[:rank]), sum(totals)
            else:
                col.get_collisions(cq, ct, rt, n_buf, rows, cap).synchronize()
                assert not col.overflowed
                local, base, total = col.local_pairs, col.pair_base, col.total_pairs
            assert (int(n_buf.item()) & 0xFFFFFFFF) == (total & 0xFFFFFFFF)
            got = rows.cpu().numpy().view(np.uint32).reshape(-1, 2)
            assert (got[local:] == 0xFFFFFFFF).all()
            np.save(os.path.join(tmpdir, "rows{}_{}.npy".format(f, rank)), got[:local])
            log.append([local, base, total, col.q_begin, col.q_end])
        if graph:
            assert col.graph_replays >= frames - 2, col.graph_replays
        np.save(os.path.join(tmpdir, "log{}.npy".format(rank)), np.array(log))
        col.close()
    finally:
        dist.destroy_process_group()


def _run(tmp_path, oracle, n, kind, mode, balance=False, graph=False, frames=1, raw=False, world=2, one_device=False):
    import torch
    import torch.multiprocessing as mp
    if not torch.cuda.is_available() or (not one_device and torch.cuda.device_count() < world):
        pytest.skip("needs {} GPUs".format(world))
    from collision_b200.sharded import ShardedCollider
    if mode.partition("/")[0] not in ShardedCollider.MODES:
        pytest.skip("mode {} not built".format(mode))
    mp.spawn(_worker, args=(world, _free_port(), n, kind, mode, balance, graph, frames, raw, str(tmp_path), one_device),
             nprocs=world, join=True)
    c, r = _scene(n, kind)
    want = oracle.collide(c, r)
    logs = [np.load(tmp_path / "log{}.npy".format(k)) for k in range(world)]
    for f in range(frames):
        parts = [np.load(tmp_path / "rows{}_{}.npy".format(f, k)) for k in range(world)]
        for k in range(world):                                                # bases = running totals, ranges tile [0, n)
            assert logs[k][f][1] == sum(logs[j][f][0] for j in range(k))
            assert logs[k][f][3] == (logs[k - 1][f][4] if k else 0)
        assert logs[-1][f][4] == n
        assert all(l[f][2] == want["total"] for l in logs)
        np.testing.assert_equal(np.concatenate(parts), want["pairs"])        # rank-major == single rank, in ORDER
    return logs


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("n,kind", [(20011, "uniform"), (16384, "duplicates")])
def test_two_ranks_equal_single_rank(tmp_path, oracle, mode, n, kind):
    _run(tmp_path, oracle, n, kind, mode)


@pytest.mark.parametrize("mode", ["sharded_build", "sharded_build/local_sort", "sharded_build/replicated_sort"])
def test_two_ranks_many_windows(tmp_path, oracle, mode):
    """Enough spheres for many windows, several tiles of the collect pass and unequal buckets per rank."""
    _run(tmp_path, oracle, 300007, "skewed", mode)


def test_two_ranks_root_exchange_too_small(tmp_path, oracle):
    _run(tmp_path, oracle, 20011, "uniform", "sharded_build/tiny_send")


@pytest.mark.parametrize("mode", ["gather_inputs", "sharded_build"])
def test_two_ranks_raw_byte_buffers(tmp_path, oracle, mode):
    _run(tmp_path, oracle, 8192, "uniform", mode, raw=True)


@pytest.mark.parametrize("mode", ["replicate_build", "gather_inputs"])   # (sharded_build cuts the queries where it cuts the windows)
def test_two_ranks_balanced(tmp_path, oracle, mode):
    logs = _run(tmp_path, oracle, 30000, "skewed", mode, balance=True, frames=6)
    first = max(logs[0][0][0], logs[1][0][0]) / (logs[0][0][2] / 2)
    last = max(logs[0][-1][0], logs[1][-1][0]) / (logs[0][-1][2] / 2)
    assert first > 1.3 and last < 1.4 and last < first - 0.3, (first, last)   # the re-cut evens the rows out
    assert logs[0][-1][4] != 15000                          # ... by moving the cut


@pytest.mark.parametrize("mode", MODES)
def test_two_ranks_graph_replay(tmp_path, oracle, mode):
    _run(tmp_path, oracle, 20011, "uniform", mode, graph=True, frames=4)


# ---- the same ranks, all on ONE GPU (gloo for the collectives, CUDA IPC between the processes): runs on a one-GPU
# box, and with more ranks than the 2-GPU tests above -- middle ranks, three owners to fetch from
@pytest.mark.parametrize("mode", MODES)
def test_three_ranks_on_one_device(tmp_path, oracle, mode):
    _run(tmp_path, oracle, 60013, "skewed", mode, world=3, one_device=True)


@pytest.mark.parametrize("mode", [m for m in MODES if m.startswith("sharded_build")])
def test_four_ranks_on_one_device(tmp_path, oracle, mode):
    _run(tmp_path, oracle, 60013, "skewed", mode, world=4, one_device=True)


@pytest.mark.parametrize("mode", ["sharded_build", "sharded_build/local_sort"])
def test_five_ranks_on_one_device_duplicates(tmp_path, oracle, mode):
    _run(tmp_path, oracle, 16384, "duplicates", mode, world=5, one_device=True)


def test_four_ranks_on_one_device_root_exchange_too_small(tmp_path, oracle):
    _run(tmp_path, oracle, 60013, "uniform", "sharded_build/tiny_send", world=4, one_device=True)
