"""world_size-2 gloo tests (CPU) of the multi-GPU plumbing: query sharding by
contiguous Morton-sorted ranges, broadcast of the traversal data, all-gather of the
per-rank row totals.  The per-range traversal itself is played by the CPU oracle here;
the CUDA kernels over sub-ranges are covered by tests/test_collision_gpu.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import golden_data as G


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, tmpdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        from collision_b200.sharded import broadcast_buffers, exchange_totals, shard_range

        # rank 0 owns the scene and builds; the others receive the read-only tree
        if rank == 0:
            c, r = G.random_spheres(n)
            res = oracle.collide(c, r)
            nodes = torch.from_numpy(res['nodes'].view(np.uint8).copy())
            boxes = torch.from_numpy(res['bounds'].view(np.uint8).reshape(-1).copy())
        else:
            nodes = torch.zeros((2 * n - 1) * 16, dtype=torch.uint8)
            boxes = torch.zeros((2 * n - 1) * 32, dtype=torch.uint8)
        broadcast_buffers([nodes, boxes], src=0)
        nodes_np = nodes.numpy().view(oracle.Node)
        boxes_np = boxes.numpy().view(np.float32).reshape(2 * n - 1, 2, 4)

        qb, qe = shard_range(n, world, rank)
        total, pairs, _ = oracle.traverse(nodes_np, boxes_np, q_begin=qb, q_end=qe)
        totals = exchange_totals(torch.tensor([total], dtype=torch.int64))
        base = int(totals[:rank].sum())
        np.save(os.path.join(tmpdir, "pairs{}.npy".format(rank)), pairs)
        np.save(os.path.join(tmpdir, "meta{}.npy".format(rank)), np.array([qb, qe, total, base, int(totals.sum())]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [1001, 4096])
def test_two_rank_sharding_matches_single_rank(oracle, tmp_path, n):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, join=True)
    c, r = G.random_spheres(n)
    want = oracle.collide(c, r)
    metas = [np.load(tmp_path / "meta{}.npy".format(k)) for k in range(world)]
    parts = [np.load(tmp_path / "pairs{}.npy".format(k)) for k in range(world)]
    # ranges tile [0, n) without gaps; bases are the running totals; global count agrees
    assert metas[0][0] == 0 and metas[-1][1] == n and metas[0][1] == metas[1][0]
    assert metas[0][3] == 0 and metas[1][3] == metas[0][2]
    assert all(m[4] == want['total'] for m in metas)
    # rank-major concatenation IS the single-rank (sequential reference) list
    np.testing.assert_equal(np.concatenate(parts), want['pairs'])


def test_shard_range_properties():
    from collision_b200.sharded import shard_range
    for n in (1, 7, 1000, 1 << 20, (1 << 27) + 3):
        for world in (1, 2, 4, 8):
            edges = [shard_range(n, world, r) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))
            sizes = [e - b for b, e in edges]
            assert max(sizes) - min(sizes) <= 1


def _gather_worker(rank, world, port, n, tmpdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from collision_b200.sharded import gather_slices, shard_range
        c, r = G.random_spheres(n)
        b, e = shard_range(n, world, rank)
        # every rank fills only its own spheres; the rest is poison
        coords = torch.full((n, 4), float("nan"), dtype=torch.float32)
        radii = torch.full((n,), float("nan"), dtype=torch.float32)
        coords[b:e] = torch.from_numpy(c[b:e])
        radii[b:e] = torch.from_numpy(r[b:e].astype(np.float32))
        gather_slices(coords, 4, n)
        gather_slices(radii, 1, n)
        np.save(os.path.join(tmpdir, "coords{}.npy".format(rank)), coords.numpy())
        np.save(os.path.join(tmpdir, "radii{}.npy".format(rank)), radii.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [1001, 4096])   # ragged shards (one broadcast per owner) and equal ones (all-gather)
def test_two_rank_input_gather(tmp_path, n):
    """mode "gather_inputs": the scene arrives sharded by sphere index and is all-gathered in place."""
    world = 2
    mp.spawn(_gather_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, join=True)
    c, r = G.random_spheres(n)
    for k in range(world):
        np.testing.assert_equal(np.load(tmp_path / "coords{}.npy".format(k)), c.astype(np.float32))
        np.testing.assert_equal(np.load(tmp_path / "radii{}.npy".format(k)), r.astype(np.float32))
