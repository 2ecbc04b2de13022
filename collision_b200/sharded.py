"""Traversal sharded over the GPUs of one box (one process per GPU, torch.distributed).

The reference is single-device.  This is the multi-GPU form BASELINE.json's
north_star prescribes: the scene is sorted, built and refitted ONCE (rank 0), the
read-only traversal data is broadcast over NVLink (NCCL), every rank walks a
contiguous range of the Morton-sorted queries and keeps its rows, and the per-rank
row totals are all-gathered so every rank knows the global count and its base
offset.  Concatenating the ranks' lists in rank order gives exactly the 1-GPU
(= sequential reference) list.

What travels (per sphere): the traversal record 32 B + the leaf box 32 B + the
sorted id 4 B = 68 B (f32); the Node array stays on rank 0.

The distributed plumbing (`shard_range`, `exchange_totals`, `broadcast_buffers`)
is device-agnostic so that it is covered by world_size-2 gloo tests on CPU.
"""
import math

import torch
import torch.distributed as dist
from numpy import dtype

from . import _lib
from .collision import MORTON_BITS, Collider


def shard_range(n, world, rank):
    """Contiguous range of sorted query positions owned by `rank`: [n*r/k, n*(r+1)/k)."""
    return (n * rank) // world, (n * (rank + 1)) // world


def rebalance_edges(edges, totals, query_weight=1):
    """Query-range edges for the NEXT call from the LAST call's per-rank row totals (SURVEY 8(e):
    "for clustered data repartition by equal pair counts using the gathered counts").

    `edges`: the world+1 sorted positions the last call was split at (edges[0] = 0, edges[-1] = n);
    `totals`: the rows each rank reported over its range.  The cost of a range is modelled as
    rows + query_weight * queries, spread evenly inside each old range (piecewise-constant
    density), and the new edges cut that cost into `world` equal parts.  Pure integer
    arithmetic on values every rank holds after the all-gather, so all ranks derive the same
    edges without another exchange; repeated over the frames of a moving scene it converges
    on the balanced split.  The row list is unaffected: rank-major concatenation of
    contiguous sorted ranges is the 1-GPU list for ANY edges."""
    world = len(totals)
    if len(edges) != world + 1 or any(b > e for b, e in zip(edges, edges[1:])):
        raise ValueError("edges must be world + 1 non-decreasing positions")
    lens = [e - b for b, e in zip(edges, edges[1:])]
    cost = [int(t) + query_weight * l for t, l in zip(totals, lens)]
    whole = sum(cost)
    if whole == 0:
        return list(edges)
    new = [edges[0]]
    seg, before = 0, 0                       # current old range and the cost of the ranges before it
    for j in range(1, world):
        # smallest position whose cost prefix reaches j/world of the whole:  before + cost[seg]*x/lens[seg] >= whole*j/world
        while seg < world - 1 and (before + cost[seg]) * world <= whole * j:
            before += cost[seg]
            seg += 1
        if cost[seg] == 0:
            x = 0
        else:
            need = whole * j - before * world            # = (target - before) * world, >= 0
            x = -((-need * lens[seg]) // (cost[seg] * world))    # ceil
        new.append(max(new[-1], min(edges[seg] + x, edges[seg + 1])))
    new.append(edges[-1])
    return new


class DensityModel:
    """Cost density over the sorted positions, refined by every (cut, per-rank cost) observation.

    `rebalance_edges` knows one number per rank and forgets it after the call, so on a scene whose
    density varies inside a rank's range the cut overshoots and wobbles by a few per cent (measured on
    4 GPUs).  This keeps a histogram of `bins` equal position ranges instead: an observation spreads each
    rank's cost over the bins its range covers in proportion to what the model already holds there
    (iterative proportional fitting), so cuts at different places add up to a finer picture, and the
    next cut is read off the histogram's prefix sums.  Plain Python float arithmetic on values every
    rank holds (the all-gathered totals): all ranks derive the same cut, no extra exchange."""

    def __init__(self, n, bins=256):
        self.n = n
        self.bins = max(1, min(bins, n))
        self.edges = [n * i // self.bins for i in range(self.bins + 1)]
        self.density = [1.0] * self.bins              # cost per position inside each bin

    def update(self, cut, cost):
        """cut: world + 1 non-decreasing positions from 0 to n; cost[r]: what range r cost."""
        if len(cut) != len(cost) + 1 or cut[0] != 0 or cut[-1] != self.n or any(a > b for a, b in zip(cut, cut[1:])):
            raise ValueError("cut must be world + 1 non-decreasing positions from 0 to n")
        e, d = self.edges, self.density
        mass = [0.0] * self.bins
        k0 = 0
        for r, c in enumerate(cost):
            lo, hi = cut[r], cut[r + 1]
            if hi <= lo:
                continue
            while e[k0 + 1] <= lo:
                k0 += 1
            parts, held, k = [], 0.0, k0
            while k < self.bins and e[k] < hi:
                o = min(hi, e[k + 1]) - max(lo, e[k])
                parts.append((k, o))
                held += d[k] * o
                k += 1
            for k, o in parts:
                mass[k] += c * (d[k] * o / held if held > 0 else o / (hi - lo))
        for k in range(self.bins):
            self.density[k] = mass[k] / (e[k + 1] - e[k])

    def cut(self, world):
        """world + 1 positions that split the modelled cost into equal parts."""
        e, d = self.edges, self.density
        mass = [d[k] * (e[k + 1] - e[k]) for k in range(self.bins)]
        whole = sum(mass)
        if not whole > 0:
            return [shard_range(self.n, world, r)[0] for r in range(world)] + [self.n]
        out, k, before = [0], 0, 0.0
        for j in range(1, world):
            target = whole * j / world
            while k < self.bins - 1 and before + mass[k] <= target:
                before += mass[k]
                k += 1
            x = math.ceil((target - before) / d[k]) if d[k] > 0 else 0
            out.append(max(out[-1], min(e[k] + x, e[k + 1])))
        out.append(self.n)
        return out


def exchange_totals(local_total, group=None):
    """all_gather of the per-rank row counts.  `local_total`: 1-element int64 tensor on
    the group's device.  Returns the (world,) int64 tensor of totals (same device);
    rank r's base offset is totals[:r].sum() and the global count totals.sum()."""
    world = dist.get_world_size(group)
    out = torch.empty(world, dtype=torch.int64, device=local_total.device)
    dist.all_gather_into_tensor(out, local_total.reshape(1), group=group)
    return out


def broadcast_buffers(buffers, src=0, group=None):
    """Broadcast each (uint8) buffer from `src`; over NCCL this rides NVLink / NVSwitch."""
    for b in buffers:
        dist.broadcast(b, src=src, group=group)


def gather_slices(buf, per_item, n, group=None):
    """In-place all-gather of a scene array: `buf` holds n items of `per_item` BYTES each and
    rank r has filled items [n*r/k, n*(r+1)/k) (shard_range); afterwards every rank holds all n.
    `buf` may be a typed tensor or raw bytes (like every other entry point): it is viewed as
    bytes, and must hold exactly n * per_item of them."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if not buf.is_contiguous():
        raise ValueError("buffer must be contiguous")
    flat = buf.view(torch.uint8).reshape(-1)
    if flat.numel() != n * per_item:
        raise ValueError("buffer holds {} bytes, expected {} items x {} bytes".format(flat.numel(), n, per_item))
    if n % world == 0:
        m = n // world * per_item
        mine = flat[rank * m:(rank + 1) * m]
        # NCCL gathers in place when the input is the rank's own piece of the output; other
        # backends get a private copy of the piece
        dist.all_gather_into_tensor(flat[:n * per_item], mine if flat.is_cuda else mine.clone(), group=group)
    else:   # ragged shards: one broadcast per owner
        for r in range(world):
            b, e = shard_range(n, world, r)
            if e > b:
                dist.broadcast(flat[b * per_item:e * per_item], src=r, group=group)


class ShardedCollider:
    """Collider.get_collisions across `world` GPUs.  Same call signature; rank `src`
    passes the scene (`coords_buf`, `radii_buf`); in mode "broadcast_tree" the other
    ranks may pass None, in mode "replicate_build" they pass buffers to receive it; in mode
    "gather_inputs" every rank passes full-size buffers of which it has filled its own slice
    of the spheres (`shard_range(size, world, rank)` in input order).

    After a call: `n_collisions_buf[0]` = global row count (low 32 bits, like the
    reference's counter) on every rank, `collisions_buf` = this rank's rows
    (`local_pairs` of them, global row index `pair_base` + i)."""

    MODES = ("broadcast_tree", "replicate_build", "gather_inputs", "sharded_build")
    WINDOW = 256      # sorted leaves per window of the build (csrc/refit.cu: BR_WINDOW)

    def __init__(self, ctx, size, ngroups, group_size, coord_dtype=dtype('float32'), group=None, src=0,
                 mode="broadcast_tree", balance=False, query_weight=1, graph=False, sharded_sort=None):
        """mode "broadcast_tree" (north_star): rank `src` sorts, builds and refits; the
        traversal data (68 B per sphere) is broadcast.
        mode "replicate_build": the INPUT (coords + radii, 20 B per sphere) is broadcast
        from `src` and every rank builds the identical tree itself -- the build is
        deterministic and bit-exact, so the result is the same; it trades redundant
        compute for 3.4x less NVLink traffic and no serial rank-0 phase.
        mode "gather_inputs": as "replicate_build", but the scene arrives sharded -- every rank
        owns the spheres [size*r/k, size*(r+1)/k) -- and is all-gathered in place (the same
        bytes over NVLink, but k host links feed it instead of one).
        mode "sharded_build": the scene arrives sharded as in "gather_inputs" and the sort is replicated, but
        the BUILD is dealt out: the 256-leaf windows of the sorted order are split between the ranks, a rank
        builds and refits its own windows only, the few subtree roots per window that reach beyond it are
        all-gathered and every rank finishes the nodes that span windows itself; the records and leaf boxes
        of other ranks' windows are then fetched ON DEMAND -- only the subtrees the rank's own queries can
        enter -- straight from the owners' HBM over NVLink (peer loads from buffers mapped with CUDA IPC,
        `collision_b200.shared`).  Query ranges = the window ranges (no `balance`).  Constructing the object
        is collective (the shared buffers are exchanged).  The SORT is dealt out too (sharded_sort):
          None (default): "exchange" on more than two ranks, "replicated" on one or two (with a single peer the
            all-gathers are cheap and hide behind the sort: measured 0.9 ms better at 2 GPUs, 3.6 ms worse at 8).
          True / "exchange": no rank ever holds all the codes or the whole scene.  The top 8-bit digit
            of the rank's own codes is counted, the counts (1 KB) are all-gathered, every rank derives the same
            plan -- which buckets cover which rank's windows -- and PUSHES the {code, id, coordinates, radius}
            of its own spheres in those buckets straight into the receiving rank's HBM (peer stores over
            NVLink, compacted in input order: csrc/sort.cu col_shard_push); a rank then sorts what it received.
            28 bytes per sphere cross NVLink once, instead of all-gathers of 4 + 20 bytes per sphere TO EVERY RANK.
          "local": every rank still gathers all the codes and the scene, but sorts only the buckets covering
            its own windows (`col_shard_sort`: count, select, collect, sort).
          False / "replicated": every rank sorts everything.
        The sorted positions are bit-identical in all three; the codes the spanning nodes need travel with the
        handed-over roots and the ids of fetched leaves with the halo.
        balance=True: after every call (with sync_totals) that WRITES rows the query ranges of
        the next call are re-cut from the all-gathered row totals (cost of a range = rows +
        query_weight * queries, accumulated in a `DensityModel`), so that a clustered scene
        called frame after frame converges on equal work per rank; the first call uses the
        equal-queries split.  Count-only calls (the sizing call of the reference's
        count-then-allocate protocol) never re-cut: the `local_pairs` they report describes the
        range the next call walks.  Because a re-cut moves rows between ranks from one frame to
        the next, size `collisions_buf` with headroom (see `suggested_capacity`) and check
        `overflowed` after a writing call.  The rows and their order are the same for any split.
        graph=True: calls with sync_totals=False replay the whole step -- kernels AND NCCL
        collectives -- as one CUDA graph from the second call with the same buffers on."""
        if mode not in self.MODES:
            raise ValueError("mode must be one of {}".format(self.MODES))
        self.mode = mode
        self.collider = Collider(ctx, size, ngroups, group_size, coord_dtype)
        self.group = group
        self.src = src
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.size = size
        self.balance = balance
        self.query_weight = query_weight
        self._density = DensityModel(size) if balance else None
        self.edges = [shard_range(size, self.world, r)[0] for r in range(self.world)] + [size]
        self.q_begin, self.q_end = self.edges[self.rank], self.edges[self.rank + 1]
        self.local_pairs = 0
        self.pair_base = 0
        self.total_pairs = 0
        self.overflowed = False       # last writing call (with sync_totals): local_pairs > n_collisions
        self._totals = None
        self.graph = bool(graph)
        self._graphs = {}
        self._graph_candidate = None
        self.graph_replays = 0
        auto = "exchange" if dist.get_world_size(group) > 2 else "replicated"
        sort_mode = {True: "exchange", False: "replicated", None: auto}.get(sharded_sort, sharded_sort)
        if sort_mode not in ("exchange", "local", "replicated"):
            raise ValueError("sharded_sort must be True / 'exchange', 'local' or False / 'replicated'")
        self.sort_mode = sort_mode if mode == "sharded_build" else "replicated"
        self.sharded_sort = self.sort_mode != "replicated"
        if mode == "sharded_build":
            if balance:
                raise ValueError("mode 'sharded_build' splits the queries where it splits the windows: no balance")
            self._setup_sharded_build(ctx)

    def _setup_sharded_build(self, ctx):
        from .shared import SharedBuffer
        c, n, k = self.collider, self.size, self.world
        L = _lib.lib()
        windows = -(-n // self.WINDOW)
        self.win_edges = [windows * r // k for r in range(k + 1)]
        self.edges = [min(w * self.WINDOW, n) for w in self.win_edges]
        self.q_begin, self.q_end = self.edges[self.rank], self.edges[self.rank + 1]
        c._ensure_scratch(None)
        isz = c.program.coord_dtype.itemsize
        # the two arrays the walk reads live in buffers every rank of the box can read in place
        self._sb_wide = SharedBuffer(ctx, c._wide_buf.numel(), self.group)
        self._sb_bounds = SharedBuffer(ctx, c._bounds_buf.numel(), self.group)
        c._wide_buf, c._bounds_buf = self._sb_wide.local, self._sb_bounds.local
        self._sb_ids = None
        if self.sharded_sort:
            # the sorted ids, too: a fetched leaf's id is the owner's to give
            self._sb_ids = SharedBuffer(ctx, c._ids_bufs[1].numel(), self.group)
            c._ids_bufs[1] = self._sb_ids.local
        self._sb_recv = []
        if self.sort_mode == "local":
            self._shard_sort_scratch = _lib.alloc_bytes(ctx, L.col_shard_sort_scratch_bytes(n))
            self._sort_range = torch.zeros(4, dtype=torch.int32, device=ctx)   # {offset, count, first, last bucket}
        if self.sort_mode == "exchange":
            import ctypes as C
            b, e = shard_range(n, k, self.rank)
            # receive buffers (codes, ids, coordinates, radii), each sized for the worst case: one bucket holds
            # the whole scene and every rank receives all of it
            self._sb_recv = [SharedBuffer(ctx, n * w, self.group) for w in (4, 4, 4 * isz, isz)]
            self._recv_tables = [(C.c_uint64 * k)(*sb.peer_ptrs) for sb in self._sb_recv]
            self._hist_own = torch.zeros(256, dtype=torch.int32, device=ctx)
            self._hist_all = torch.zeros(256 * k, dtype=torch.int32, device=ctx)
            self._plan = torch.zeros(8 * k, dtype=torch.int32, device=ctx)
            pos = []
            for qb, qe in zip(self.edges, self.edges[1:]):      # own windows plus one position either side
                pos += [max(qb, 1) - 1, min(qe, n - 1)] if qe > qb else [1, 0]
            self._pos_table = torch.tensor(pos, dtype=torch.int32, device=ctx)
            self._push_scratch = _lib.alloc_bytes(ctx, L.col_shard_push_scratch_bytes(e - b, k))
            self._recv_sort_scratch = _lib.alloc_bytes(ctx, L.col_shard_sort_received_scratch_bytes(n))
            self._token = torch.zeros(1, dtype=torch.int32, device=ctx)
        most = max(e - b for b, e in zip(self.edges, self.edges[1:]))
        self._cap = max(1024, most // 16)                    # subtree roots a rank may hand over (typically 1.6 % of its leaves)
        self._list_cap = max(4096, n // 16)                  # remote subtrees a rank may ask for
        # the roots the rank's windows hand over: its probe list (stays here)
        self._packed = _lib.alloc_bytes(ctx, L.col_shard_packed_bytes(self._cap, c.program._code))
        # what travels: the rank's few range roots + the indices of the spanning nodes it built (as many as roots,
        # less a few).  After every call that syncs, the exchange is cut down to 1.5 x the most any rank listed
        # (+ 1024) -- the all-gather moves whole buffers, used or not.  Identical on every rank (derived from the
        # gathered headers).
        nb = L.col_shard_exchange_bytes(self._cap, c.program._code)
        self._xchg = _lib.alloc_bytes(ctx, nb)
        self._gathered = _lib.alloc_bytes(ctx, nb * k)
        self._send_cap = self._cap
        self.sort_count = n           # pairs this rank sorted in the last call that synced
        self._list = _lib.alloc_bytes(ctx, L.col_shard_list_bytes(self._list_cap))
        self._win_edges_dev = torch.tensor(self.win_edges, dtype=torch.int32, device=ctx)
        self._side = torch.cuda.Stream(_lib.device_of(ctx))
        dist.barrier(group=self.group)

    def _packed_bytes(self):
        return _lib.lib().col_shard_exchange_bytes(self._send_cap, self.collider.program._code)

    def _check_shard_overflow(self):
        """After a call that synced.  Returns True when the call has to be repeated (the cut-down exchange was
        too small for this frame: it is reset to the full buffers first)."""
        hdr = self._gathered.view(torch.int32)
        stride = self._packed_bytes() // 4
        heads = torch.stack([hdr[r * stride:r * stride + 4] for r in range(self.world)]).tolist()
        over_roots = any(h[1] for h in heads)      # a probe list or the range roots: fixed sizes
        over_span = any(h[3] for h in heads)
        over_list = bool(int(self._list.view(torch.int32)[1].item()))
        if over_span and not over_roots and self._send_cap < self._cap:
            self._send_cap = self._cap
            return True
        if over_roots or over_span or over_list:
            raise RuntimeError("sharded_build: more subtree roots / remote subtrees than the exchange buffers hold "
                               "(cap {}, list {}); use mode 'gather_inputs' for this scene".format(self._cap, self._list_cap))
        most = max(h[2] for h in heads)
        self._send_cap = min(self._cap, max(self._send_cap if self._send_cap < self._cap else 0, most + most // 2 + 1024))
        if self.sort_mode == "exchange":
            self.sort_count = int(self._plan[8 * self.rank + 1].item())
        elif self.sort_mode == "local":
            self.sort_count = int(self._sort_range[1].item())
        return False

    def close(self):
        """Collective teardown: drops the captured graphs (they hold NCCL work) and unmaps / frees the
        buffers shared between the ranks.  Call it on every rank before destroy_process_group()."""
        self._graphs.clear()
        self._graph_candidate = None
        torch.cuda.synchronize()
        if self.mode == "sharded_build" and getattr(self, "_sb_wide", None) is not None:
            c = self.collider
            c._wide_buf = c._bounds_buf = None
            c._bufs_for = None
            self._sb_wide.close()
            self._sb_bounds.close()
            self._sb_wide = self._sb_bounds = None
            if self._sb_ids is not None:
                c._ids_bufs[1] = None
                self._sb_ids.close()
                self._sb_ids = None
            for sb in self._sb_recv:
                sb.close()
            self._sb_recv = []

    def suggested_capacity(self, headroom=1.5):
        """Rows to make room for on this rank after a count-only call: with balance=True the
        next frames move the cuts, so a rank may receive up to its fair share of the whole list
        plus the drift of the density model -- `headroom` x max(local, total / world)."""
        fair = -(-self.total_pairs // self.world)
        if not self.balance:
            return self.local_pairs
        return int(headroom * max(self.local_pairs, fair)) + 1

    # instrumentation is the inner collider's
    @property
    def stage_events(self):
        return self.collider.stage_events

    @stage_events.setter
    def stage_events(self, v):
        self.collider.stage_events = v

    @property
    def kernel_launches(self):
        return self.collider.kernel_launches

    def _recut(self, totals):
        """New query ranges from the rows every rank just reported (identical on every rank: same
        edges, same gathered totals, same arithmetic)."""
        cost = [rows + self.query_weight * (e - b) for rows, b, e in zip(totals, self.edges, self.edges[1:])]
        self._density.update(self.edges, cost)
        self.edges = self._density.cut(self.world)
        self.q_begin, self.q_end = self.edges[self.rank], self.edges[self.rank + 1]

    def _shared_buffers(self):
        c = self.collider
        n = c.size
        isz = c.program.coord_dtype.itemsize
        leaf_boxes = c._bounds_buf[(n - 1) * 8 * isz:(2 * n - 1) * 8 * isz]
        return [c._wide_buf, leaf_boxes, c._ids_bufs[1][:n * 4], c._scene_buf]

    def get_collisions(self, cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf, n_collisions,
                       wait_for=None, sync_totals=True):
        """n_collisions: capacity (rows) of THIS rank's collisions_buf."""
        if collisions_buf is None and n_collisions > 0:
            raise ValueError("Invalid collisions_buf for n_collisions > 0")
        c = self.collider
        c._ensure_scratch(cq)
        _lib.begin(cq, wait_for)
        use_graph = self.graph and not sync_totals and c.stage_events is None
        if use_graph:
            key = (cq.cuda_stream, self.size, n_collisions, self.q_begin, self.q_end, _lib.ptr(coords_buf),
                   _lib.ptr(radii_buf), _lib.ptr(n_collisions_buf), _lib.ptr(collisions_buf), c._trv_scratch_bytes,
                   getattr(self, "_send_cap", 0))
            entry = self._graphs.get(key)
            if entry is None and self._graph_candidate == key:
                # second call with this signature (scratch allocated, NCCL communicator warm): capture
                g = torch.cuda.CUDAGraph()
                before = c.kernel_launches
                with torch.cuda.graph(g, stream=cq, capture_error_mode="thread_local"):
                    self._enqueue(cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf, n_collisions)
                entry = self._graphs[key] = (g, c.kernel_launches - before, self._totals)
                c.kernel_launches = before
                while len(self._graphs) > Collider.MAX_GRAPHS:
                    self._graphs.pop(next(iter(self._graphs)))
            if entry is not None:
                with torch.cuda.stream(cq):
                    entry[0].replay()
                c.kernel_launches += entry[1]
                self._totals = entry[2]
                self.graph_replays += 1
                return _lib.done(cq)
            self._graph_candidate = key
        self._enqueue(cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf, n_collisions)
        ev = _lib.done(cq)
        if sync_totals:     # host-visible bookkeeping (one small readback)
            ev.synchronize()
            t = self._totals.tolist()
            self.local_pairs = t[self.rank]
            self.pair_base = sum(t[:self.rank])
            self.total_pairs = sum(t)
            if self.mode == "sharded_build" and self._check_shard_overflow():
                # (every rank takes this branch together: the decision comes from the gathered headers)
                return self.get_collisions(cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf, n_collisions,
                                           None, True)
            if n_collisions > 0:
                self.overflowed = self.local_pairs > n_collisions
                if self.balance:     # (count-only calls never move the cuts: see __init__)
                    self._recut(t)
        return ev

    def _front_sharded(self, cq, coords_buf, radii_buf):
        """Scene bounds and Morton codes from the rank's OWN slice of the spheres (a 32-byte all-reduce and
        an all-gather of 4 bytes per sphere), then the replicated sort -- while the all-gather of the
        coordinates and radii themselves (20 bytes per sphere, needed only by the leaf boxes of the build)
        runs beside it on a second stream."""
        c = self.collider
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr(cq)
        n, code, isz = self.size, c.program._code, c.program.coord_dtype.itemsize
        b, e = shard_range(n, self.world, self.rank)
        fdt = torch.float32 if isz == 4 else torch.float64
        scene = c._scene_buf.view(fdt)                       # {min[4], max[4]}
        if e > b:
            _lib.check(L.col_bounds(P(coords_buf) + b * 4 * isz, e - b, code, 4, P(c._scene_buf),
                                    P(c.reducer._ensure_scratch(cq)), S), "col_bounds")
        else:
            scene[:4] = float("inf")
            scene[4:] = float("-inf")
        # one MIN all-reduce of {min, -max}: exact (negation and min are), so every rank holds the bounds a
        # single device would have computed
        scene[4:].neg_()
        dist.all_reduce(scene, op=dist.ReduceOp.MIN, group=self.group)
        scene[4:].neg_()
        c._mark(cq, "bounds", 1)
        if self.sort_mode == "exchange":
            return self._front_exchange(cq, coords_buf, radii_buf, b, e)
        if e > b:
            _lib.check(L.col_morton(P(c._codes_bufs[0]) + 4 * b, None, P(coords_buf) + b * 4 * isz, P(c._scene_buf),
                                    code, e - b, e - b, S), "col_morton")
        gather_slices(c._codes_bufs[0][:4 * n], 4, n, self.group)
        if self.sort_mode == "replicated":
            _lib.check(L.col_range(P(c._ids_bufs[0]), n, S), "col_range")
        c._mark(cq, "morton", 1 if self.sharded_sort else 2)
        # the scene itself travels on the side stream, behind the codes in NCCL's order
        ready = torch.cuda.Event()
        ready.record(cq)
        with torch.cuda.stream(self._side):
            self._side.wait_event(ready)
            gather_slices(coords_buf, 4 * isz, n, self.group)
            gather_slices(radii_buf, isz, n, self.group)
            gathered = torch.cuda.Event()
            gathered.record(self._side)
        if self.sort_mode == "local":
            # own windows plus one position either side (the window pass reads its neighbours' codes)
            lo, hi = max(self.q_begin, 1) - 1, min(self.q_end, n - 1)
            if self.q_end > self.q_begin:
                _lib.check(L.col_shard_sort(P(c._codes_bufs[0]), n, lo, hi, MORTON_BITS, P(c._codes_bufs[1]),
                                            P(c._ids_bufs[1]), P(self._sort_range), P(self._shard_sort_scratch), S),
                           "col_shard_sort")
            c._mark(cq, "sort", 5 + (MORTON_BITS + 7) // 8 + 2)   # count, bases, select, collect; histogram, bases, passes
        else:
            c._sort_codes(cq)
        cq.wait_event(gathered)
        c._mark(cq, "broadcast", 0)      # (what is left of the scene's all-gather after the sort)

    def _front_exchange(self, cq, coords_buf, radii_buf, b, e):
        """Own codes -> top-digit counts (all-gathered: 1 KB per rank) -> push of the own spheres to the ranks
        whose windows they fall into -> sort of what this rank received."""
        c = self.collider
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr(cq)
        n, code, isz, me = self.size, c.program._code, c.program.coord_dtype.itemsize, self.rank
        own_coords, own_radii = P(coords_buf) + b * 4 * isz, P(radii_buf) + b * isz
        if e > b:
            _lib.check(L.col_morton(P(c._codes_bufs[0]), None, own_coords, P(c._scene_buf), code, e - b, e - b, S),
                       "col_morton")
        _lib.check(L.col_shard_count(P(c._codes_bufs[0]), e - b, MORTON_BITS, P(self._hist_own), S), "col_shard_count")
        dist.all_gather_into_tensor(self._hist_all, self._hist_own, group=self.group)
        c._mark(cq, "morton", 2)
        _lib.check(L.col_shard_push(P(self._hist_all), self.world, me, P(self._pos_table), n, P(c._codes_bufs[0]),
                                    e - b, b, own_coords, own_radii, code, MORTON_BITS, *self._recv_tables,
                                    P(self._plan), P(self._push_scratch), S), "col_shard_push")
        # every rank's pushes precede its part of this collective in stream order: past it, all have landed
        dist.all_reduce(self._token, group=self.group)
        c._mark(cq, "exchange", 1 + self.world)
        if self.q_end > self.q_begin:
            recv_codes, recv_ids = self._sb_recv[0].local, self._sb_recv[1].local
            _lib.check(L.col_shard_sort_received(P(recv_codes), P(recv_ids), n, P(self._plan) + 32 * me, MORTON_BITS,
                                                 P(c._codes_bufs[1]), P(c._ids_bufs[0]), P(c._ids_bufs[1]),
                                                 P(self._recv_sort_scratch), S), "col_shard_sort_received")
        c._mark(cq, "sort", 4 + (MORTON_BITS + 7) // 8)

    def _build_sharded(self, cq, coords_buf, radii_buf):
        """Own windows -> all-gather of the handed-over roots -> spanning nodes (replicated) -> halo fetch."""
        c = self.collider
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr(cq)
        n, code = self.size, c.program._code
        leaf_ids = c._ids_bufs[1]
        if self.sort_mode == "exchange":
            # the window pass finds a leaf's sphere through its PLACE in the receive buffers (the leaf records
            # of the reference-layout Node array, which nothing reads in this mode, carry that place too)
            leaf_ids, coords_buf, radii_buf = c._ids_bufs[0], self._sb_recv[2].local, self._sb_recv[3].local
        _lib.check(L.col_shard_build_windows(P(c._codes_bufs[1]), P(leaf_ids), P(c._nodes_buf), P(c._wide_buf),
                                             P(c._flags_buf), P(c._bounds_buf), P(coords_buf), P(radii_buf),
                                             P(c._scene_buf), code, n, self.win_edges[self.rank],
                                             self.win_edges[self.rank + 1], P(self._packed), self._cap, S),
                   "col_shard_build_windows")
        c._mark(cq, "build_windows", 2)
        _lib.check(L.col_shard_top_local(P(c._codes_bufs[1]), P(c._nodes_buf), P(c._wide_buf), P(c._flags_buf),
                                         P(c._bounds_buf), P(c._scene_buf), code, n, self.q_begin, self.q_end,
                                         P(self._packed), self._cap, P(self._xchg), self._send_cap, S),
                   "col_shard_top_local")
        c._mark(cq, "build_top_local", 1)
        nb = self._packed_bytes()
        dist.all_gather_into_tensor(self._gathered[:nb * self.world], self._xchg[:nb], group=self.group)
        c._mark(cq, "gather_roots", 0)
        _lib.check(L.col_shard_top(P(self._gathered), self._send_cap, self.world, self.rank, P(c._codes_bufs[1]),
                                   P(c._nodes_buf), P(c._wide_buf), P(c._flags_buf), P(c._bounds_buf),
                                   P(c._scene_buf), code, n, P(self._sb_wide.table), S), "col_shard_top")
        c._mark(cq, "build_top", 3)
        _lib.check(L.col_shard_halo(P(self._packed), self._cap, P(self._gathered), self._send_cap, self.world, self.rank,
                                    P(c._wide_buf),
                                    P(c._bounds_buf), P(c._scene_buf), code, P(c._flags_buf), P(self._list),
                                    self._list_cap, P(self._sb_wide.table), P(self._sb_bounds.table),
                                    P(self._win_edges_dev), n, self.q_begin, self.q_end,
                                    P(self._sb_ids.table) if self.sharded_sort else None,
                                    P(c._ids_bufs[1]) if self.sharded_sort else None, S), "col_shard_halo")
        c._mark(cq, "halo", 2)

    def _enqueue(self, cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf, n_collisions):
        c = self.collider
        c._mark(cq, "begin", 0)
        with torch.cuda.stream(cq):
            if self.mode == "broadcast_tree":
                if self.rank == self.src:
                    c._sort_stage(cq, coords_buf)
                    c._build_wide(cq, coords_buf, radii_buf)
                broadcast_buffers(self._shared_buffers(), self.src, self.group)
                c._mark(cq, "broadcast", 0)
            elif self.mode == "sharded_build":
                self._front_sharded(cq, coords_buf, radii_buf)
                self._build_sharded(cq, coords_buf, radii_buf)
            else:
                if self.mode == "gather_inputs":
                    isz = c.program.coord_dtype.itemsize
                    gather_slices(coords_buf, 4 * isz, self.size, self.group)
                    gather_slices(radii_buf, isz, self.size, self.group)
                else:
                    broadcast_buffers([coords_buf, radii_buf], self.src, self.group)
                c._mark(cq, "broadcast", 0)
                c._sort_stage(cq, coords_buf)
                c._build_wide(cq, coords_buf, radii_buf)
            c._walk_range(cq, collisions_buf, n_collisions, self.q_begin, self.q_end)
            self._totals = exchange_totals(c._total_buf, self.group)
            total = self._totals.sum().reshape(1)
            _lib.check(_lib.lib().col_copy(_lib.ptr(n_collisions_buf), _lib.ptr(total), 4, _lib.stream_ptr(cq)),
                       "col_copy")
            c._mark(cq, "all_gather", 0)
