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


def _gather_worker(rank, world, port, n, tmpdir, raw):
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
        if raw:     # raw-byte buffers, like every other entry point takes them
            coords, radii = coords.view(torch.uint8).reshape(-1), radii.view(torch.uint8).reshape(-1)
        gather_slices(coords, 16, n)
        gather_slices(radii, 4, n)
        np.save(os.path.join(tmpdir, "coords{}.npy".format(rank)), coords.view(torch.float32).reshape(n, 4).numpy())
        np.save(os.path.join(tmpdir, "radii{}.npy".format(rank)), radii.view(torch.float32).numpy())
        if rank == 0:
            with pytest.raises(ValueError):
                gather_slices(torch.zeros(n * 16 - 4, dtype=torch.uint8), 16, n)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("raw", [False, True], ids=["typed", "raw-bytes"])
@pytest.mark.parametrize("n", [1001, 4096])   # ragged shards (one broadcast per owner) and equal ones (all-gather)
def test_two_rank_input_gather(tmp_path, n, raw):
    """mode "gather_inputs": the scene arrives sharded by sphere index and is all-gathered in place."""
    world = 2
    mp.spawn(_gather_worker, args=(world, _free_port(), n, str(tmp_path), raw), nprocs=world, join=True)
    c, r = G.random_spheres(n)
    for k in range(world):
        np.testing.assert_equal(np.load(tmp_path / "coords{}.npy".format(k)), c.astype(np.float32))
        np.testing.assert_equal(np.load(tmp_path / "radii{}.npy".format(k)), r.astype(np.float32))


# ------------------------------------------------------------------ balanced split
def test_rebalance_edges_properties():
    from collision_b200.sharded import rebalance_edges, shard_range
    rng = np.random.RandomState(7)
    for world in (1, 2, 4, 8):
        for n in (1, 9, 1000, (1 << 27) + 3):
            edges = [shard_range(n, world, r)[0] for r in range(world)] + [n]
            for trial in range(20):
                totals = [int(t) for t in rng.randint(0, 1 << 40, size=world) * rng.randint(0, 2, size=world)]
                totals = [t if e > b else 0 for t, b, e in zip(totals, edges, edges[1:])]   # no queries, no rows
                for w in (0, 1, 16):
                    new = rebalance_edges(edges, totals, w)
                    assert len(new) == world + 1 and new[0] == 0 and new[-1] == n
                    assert all(a <= b for a, b in zip(new, new[1:]))
                    # under the piecewise-constant model every new range carries whole/world (+- one query's cost)
                    cost = [t + w * (e - b) for t, b, e in zip(totals, edges, edges[1:])]
                    whole = sum(cost)
                    if whole == 0:
                        assert new == edges
                        continue

                    def prefix(x):      # modelled cost of the positions [0, x), as an exact fraction * world
                        c = 0.0
                        for k in range(world):
                            if edges[k + 1] > edges[k]:
                                c += cost[k] * (min(max(x, edges[k]), edges[k + 1]) - edges[k]) / (edges[k + 1] - edges[k])
                        return c
                    per_query = max(cst / max(e - b, 1) for cst, b, e in zip(cost, edges, edges[1:]))
                    for j in range(1, world):
                        # the cut sits at the first position whose cost prefix reaches j/world of the whole
                        assert prefix(new[j]) >= whole * j / world * (1 - 1e-9) - 1e-6
                        assert prefix(new[j]) - whole * j / world <= per_query * (1 + 1e-9) + 1e-6
    # the same totals over unequal edges; all-zero ranges keep their positions only if nothing was seen anywhere
    assert rebalance_edges([0, 50, 100], [100, 0], 0) == [0, 25, 100]
    assert rebalance_edges([0, 25, 50, 75, 100], [0, 0, 0, 0], 0) == [0, 25, 50, 75, 100]
    assert rebalance_edges([0, 0, 0, 100], [0, 0, 50], 1) == [0, 34, 67, 100]
    with pytest.raises(ValueError):
        rebalance_edges([0, 60, 50, 100], [1, 1, 1])
    with pytest.raises(ValueError):
        rebalance_edges([0, 50, 100], [1, 1, 1])


def test_rebalance_converges_on_a_skewed_density():
    """Frame after frame on a fixed scene the per-rank rows approach whole/world."""
    from collision_b200.sharded import rebalance_edges, shard_range
    n, world = 1 << 16, 8
    rng = np.random.RandomState(3)
    rows = (rng.pareto(3.0, size=n) * 8).astype(np.int64)           # skewed rows per query ...
    rows[: n // 8] *= 20                                             # ... and one dense region
    csum = np.concatenate([[0], np.cumsum(rows)])
    edges = [shard_range(n, world, r)[0] for r in range(world)] + [n]
    spread = []
    for frame in range(6):
        totals = [int(csum[e] - csum[b]) for b, e in zip(edges, edges[1:])]
        spread.append(max(totals) / (sum(totals) / world))
        edges = rebalance_edges(edges, totals, 0)
    assert spread[0] > 3.0           # the equal-queries split is badly skewed here
    assert spread[-1] < 1.05         # the re-cut split is balanced to 5 %
    assert max(spread[2:]) < 1.05    # and stays there (it may wobble by a query or two)


def test_density_model_converges_without_wobble():
    """ShardedCollider(balance=True) keeps a histogram of the cost over the sorted positions: cuts at
    different places add up, so the split settles instead of overshooting back and forth."""
    from collision_b200.sharded import DensityModel, shard_range
    n = 1 << 20
    rng = np.random.RandomState(1)
    rows = np.ones(n, np.int64)                   # 64 clusters of different density in the first 55 %, sparse after
    span = n * 55 // 100 // 64
    for k in range(64):
        rows[k * span:(k + 1) * span] = rng.randint(5, 80)
    csum = np.concatenate([[0], np.cumsum(rows)])
    for world in (2, 4, 8):
        model = DensityModel(n)
        cut = [shard_range(n, world, r)[0] for r in range(world)] + [n]
        assert model.cut(world) == cut            # nothing seen yet: the equal-queries split
        spread = []
        for frame in range(8):
            cost = [int(csum[e] - csum[b]) + (e - b) for b, e in zip(cut, cut[1:])]
            spread.append(max(cost) * world / sum(cost))
            model.update(cut, cost)
            cut = model.cut(world)
            assert cut[0] == 0 and cut[-1] == n and all(a <= b for a, b in zip(cut, cut[1:]))
        assert spread[0] > 1.7 and spread[1] < 1.2 and max(spread[4:]) < 1.04, (world, spread)
    # degenerate inputs: empty ranges, zero cost everywhere, one position, more ranks than positions
    model = DensityModel(10)
    model.update([0, 0, 10, 10], [0, 50, 0])
    assert model.cut(2) == [0, 5, 10]
    model.update([0, 5, 10], [0, 0])
    assert model.cut(2) == [0, 5, 10]             # no cost seen: equal queries
    assert DensityModel(1).cut(4) in ([0, 0, 0, 0, 1], [0, 1, 1, 1, 1])
    with pytest.raises(ValueError):
        DensityModel(10).update([0, 6, 5, 10], [1, 1, 1])
    with pytest.raises(ValueError):
        DensityModel(10).update([0, 5, 9], [1, 1])


def test_sharded_collider_recut_bookkeeping():
    """The host-side step ShardedCollider(balance=True) takes after the totals are gathered (no GPU needed:
    the object is assembled by hand around the bookkeeping fields)."""
    from collision_b200.sharded import DensityModel, ShardedCollider, shard_range
    n, world = 1000, 4
    ranks = []
    for rank in range(world):
        sc = ShardedCollider.__new__(ShardedCollider)
        sc.size, sc.world, sc.rank, sc.query_weight = n, world, rank, 1
        sc.edges = [shard_range(n, world, r)[0] for r in range(world)] + [n]
        sc._density = DensityModel(n)
        ranks.append(sc)
    for totals in ([9000, 500, 250, 250], [3000, 3000, 2000, 2000]):
        for sc in ranks:
            sc._recut(totals)
        assert all(sc.edges == ranks[0].edges for sc in ranks)
        assert all((sc.q_begin, sc.q_end) == (sc.edges[sc.rank], sc.edges[sc.rank + 1]) for sc in ranks)
        assert ranks[0].edges[0] == 0 and ranks[0].edges[-1] == n
    assert ranks[0].edges[1] < 250           # the dense first range was cut well below the equal split


def _balance_worker(rank, world, port, n, tmpdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        from collision_b200.sharded import DensityModel, exchange_totals, shard_range
        c, r = _skewed_scene(n)
        model = DensityModel(n)
        res = oracle.collide(c, r)      # every rank builds the identical tree ("replicate_build")
        edges = [shard_range(n, world, k)[0] for k in range(world)] + [n]
        log = []
        for frame in range(3):
            total, pairs, _ = oracle.traverse(res['nodes'], res['bounds'], q_begin=edges[rank], q_end=edges[rank + 1])
            totals = exchange_totals(torch.tensor([total], dtype=torch.int64)).tolist()
            np.save(os.path.join(tmpdir, "pairs{}_{}.npy".format(frame, rank)), pairs)
            log.append(list(edges) + totals)
            model.update(edges, [t + (e - b) for t, b, e in zip(totals, edges, edges[1:])])
            edges = model.cut(world)
        np.save(os.path.join(tmpdir, "log{}.npy".format(rank)), np.array(log))
    finally:
        dist.destroy_process_group()


def _skewed_scene(n):
    """Half the spheres packed into a small corner (many overlaps), half spread out."""
    c, r = G.random_spheres(n)
    c = c.copy()
    c[: n // 2, :3] *= 0.2
    return c, r


def test_two_rank_rebalanced_split_keeps_the_list(oracle, tmp_path):
    n, world = 2000, 2
    mp.spawn(_balance_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, join=True)
    c, r = _skewed_scene(n)
    want = oracle.collide(c, r)
    logs = [np.load(tmp_path / "log{}.npy".format(k)) for k in range(world)]
    np.testing.assert_equal(logs[0], logs[1])            # every rank derived the same edges from the same totals
    for frame in range(3):
        parts = [np.load(tmp_path / "pairs{}_{}.npy".format(frame, k)) for k in range(world)]
        # any contiguous split gives the single-rank (sequential reference) list
        np.testing.assert_equal(np.concatenate(parts), want['pairs'])
        assert logs[0][frame][world + 1:].sum() == want['total']
    first, last = logs[0][0][world + 1:], logs[0][-1][world + 1:]
    assert first.max() / first.sum() > 0.6               # equal-queries split: one rank holds most rows
    assert last.max() / last.sum() < first.max() / first.sum()
    assert (logs[0][0][:world + 1] != logs[0][-1][:world + 1]).any()
