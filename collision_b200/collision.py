"""LBVH sphere-collision pipeline (mirror of collision/collision.py upstream).

Collider keeps the reference's constructor, `resize`, `n_nodes`, `padded_size`
and `get_collisions(cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf,
n_collisions, wait_for)`; Node / NO_NODE and the 16-byte Node, 32/64-byte Bound
and 4-lane coordinate layouts are unchanged.  Handles map as
ctx -> torch.device('cuda:i'), cq -> torch.cuda.Stream, cl.Buffer -> torch.Tensor,
cl.Event -> torch.cuda.Event.

Stages (all hand-written sm_100a kernels behind include/collision_b200.h):
  scene bounds (one launch)        <- bounds1 + bounds2           collision.py:157-159
  Morton codes + ids (one launch)  <- fill + range + calculateCodes        :138-165
  one-sweep stable sort            <- RadixSorter.sort, 8 x (5..9 launches) :167-169
  Karras tree (one launch)         <- fillInternal + generateBVH           :171-180
  leaf boxes + refit (one launch)  <- leafBounds + internalBounds          :181-190
  traversal                        <- traverse                              :191-196
Traversal commits pairs either in the deterministic order of the reference's
work-items run sequentially (count -> scan -> write; the default) or through one
atomic counter in arbitrary order exactly like upstream (`ordered=False`).
"""
import contextlib

import torch
from numpy import dtype, iinfo

from . import _lib
from .bounds import Bounds
from .misc import Program, roundUp, same_context
from .misc import np_float_dtypes as _float_names
from .radix import RadixSorter

np_float_dtypes = set(map(dtype, _float_names))

Node = dtype([('parent', 'uint32'), ('right_edge', 'uint32'), ('data', 'uint32', 2)])

NO_NODE = iinfo(Node.fields['parent'][0]).max

MORTON_BITS = 30   # 3 x 10 bits (collision.cl:23)


class CollisionProgram(Program):
    """collision/collision.py:13-29.  `kernels` exposes the reference's kernel names
    with their argument lists -- (cq, global_size, local_size, *args, wait_for=None)
    -- because upstream's tests launch them directly (tests/test_collision.py)."""

    def __init__(self, ctx, coord_dtype=dtype('float32')):
        coord_dtype = dtype(coord_dtype)
        if coord_dtype not in np_float_dtypes:
            raise ValueError("Invalid dtype: {}".format(coord_dtype))
        if coord_dtype == dtype('float16'):
            raise ValueError("Unsupported dtype: float16 (float32 / float64)")
        self.coord_dtype = coord_dtype
        super().__init__(ctx)
        self._code = _lib.float_code(coord_dtype)
        L = _lib.lib()
        code = self._code

        def _range(cq, gs, ls, idxs):
            _lib.check(L.col_range(_lib.ptr(idxs), gs, _lib.stream_ptr(cq)), "col_range")

        def calc_codes(cq, gs, ls, codes, coords, rng, n):
            _lib.check(L.col_morton(_lib.ptr(codes), None, _lib.ptr(coords), _lib.ptr(rng), code, n, n,
                                    _lib.stream_ptr(cq)), "col_morton")

        def fill_internal(cq, gs, ls, nodes, ids, n):
            _lib.check(L.col_fill_leaves(_lib.ptr(nodes), _lib.ptr(ids), n, _lib.stream_ptr(cq)), "col_fill_leaves")

        def generate_bvh(cq, gs, ls, codes, nodes, n):
            _lib.check(L.col_build_tree(_lib.ptr(codes), None, _lib.ptr(nodes), n, _lib.stream_ptr(cq)),
                       "col_build_tree")

        def leaf_bounds(cq, gs, ls, bounds, coords, radii, nodes, n):
            _lib.check(L.col_leaf_bounds(_lib.ptr(bounds), _lib.ptr(coords), _lib.ptr(radii), _lib.ptr(nodes),
                                         code, n, _lib.stream_ptr(cq)), "col_leaf_bounds")

        def internal_bounds(cq, gs, ls, bounds, flags, nodes, n):
            _lib.check(L.col_refit(_lib.ptr(bounds), _lib.ptr(flags), _lib.ptr(nodes), None, None, code, n,
                                   _lib.stream_ptr(cq)), "col_refit")

        def traverse(cq, gs, ls, collisions, nxt, n_collisions, nodes, bounds, n):
            _lib.check(L.col_traverse_atomic(_lib.ptr(collisions), _lib.ptr(nxt), n_collisions, _lib.ptr(nodes),
                                             _lib.ptr(bounds), code, n, 0, n, _lib.stream_ptr(cq)),
                       "col_traverse_atomic")

        kernel = _lib.kernel_launcher
        self.kernels = {
            'range': kernel(_range),
            'calculateCodes': kernel(calc_codes),
            'fillInternal': kernel(fill_internal),
            'generateBVH': kernel(generate_bvh),
            'leafBounds': kernel(leaf_bounds),
            'internalBounds': kernel(internal_bounds),
            'traverse': kernel(traverse),
        }


class Collider:
    code_dtype = dtype('uint32')
    flag_dtype = dtype('uint32')
    counter_dtype = dtype('uint32')
    id_dtype = dtype('uint32')
    MAX_GRAPHS = 8      # captured launch sequences kept per Collider (keyed on the buffers of a call)

    def __init__(self, ctx, size, ngroups, group_size, coord_dtype=dtype('float32'),
                 program=None, sorter_programs=(None, None), reducer_program=None, ordered=True,
                 layout="wide", graph=False, fused_build=True):
        coord_dtype = dtype(coord_dtype)
        self.size = size
        # True: tree + leaf boxes + refit in one bottom-up launch (col_build_refit_wide);
        # False: Karras' searches, then the climb (col_build_tree_wide + col_refit_wide) -- same arrays
        self.fused_build = fused_build
        self.group_size = group_size
        self.ordered = ordered
        # "wide": traversal records + single-pass compaction (default, needs ordered=True);
        # "reference": only the reference's Node / Bound arrays are used
        if layout not in ("wide", "reference"):
            raise ValueError("layout must be 'wide' or 'reference'")
        self.fast = layout == "wide" and ordered

        self.sorter = RadixSorter(
            ctx, self.padded_size, group_size,
            key_dtype=self.code_dtype, value_dtype=self.id_dtype,
            program=sorter_programs[0], scan_program=sorter_programs[1])
        self.reducer = Bounds(ctx, ngroups, group_size, coord_dtype=dtype((coord_dtype, 3)),
                              program=reducer_program)
        if program is None:
            program = CollisionProgram(ctx, coord_dtype)
        else:
            if not same_context(program.context, ctx):
                raise ValueError("Collider and program context must match")
            if program.coord_dtype != coord_dtype:
                raise ValueError("Collider and program coord_dtype must match")
        self.program = program
        self._bufs_for = None   # device scratch is allocated on first use
        # optional instrumentation: set `stage_events = []` to have every call
        # append (stage name, torch.cuda.Event) marks on the launching stream;
        # kernel_launches counts the kernels this object has enqueued
        self.stage_events = None
        self.kernel_launches = 0
        # graph=True: the launch sequence of get_collisions is captured into a CUDA graph the
        # second time it is called with the same buffers, sizes and stream, and replayed from
        # then on (one graph launch instead of ~17 enqueues: the short kernels at the head of
        # the pipeline are otherwise bound by the host's enqueue rate)
        self.graph = bool(graph)
        self._graphs = {}
        self._graph_candidate = None
        # measurement aid: a ctypes array of three cudaEvent_t handles recorded inside the
        # traversal call (after the walk, the confirmation, the scan); see bench.py
        self._trv_marks = None

    # -- scratch ----------------------------------------------------------
    def _ensure_scratch(self, cq=None):
        key = (self.size, self.padded_size)
        if self._bufs_for == key:
            return
        # the scratch is only ever used by work enqueued on `cq`: allocate it with `cq` current so
        # that the caching allocator re-uses a dropped block in cq's order (a block allocated on
        # another stream could be handed out again while kernels on cq still read it)
        with torch.cuda.stream(cq) if cq is not None else contextlib.nullcontext():
            self._allocate(key)

    def _allocate(self, key):
        ctx = self.program.context
        L = _lib.lib()
        self._graphs.clear()          # captured graphs point into the scratch being replaced
        self._graph_candidate = None
        # can't sort in place (collision.py:60-69)
        self._ids_bufs = [_lib.alloc_bytes(ctx, self.padded_size * 4) for _ in range(2)]
        self._codes_bufs = [_lib.alloc_bytes(ctx, self.padded_size * 4) for _ in range(2)]
        self._nodes_buf = _lib.alloc_bytes(ctx, self.n_nodes * Node.itemsize)
        # dual use: scene bounds first, per-node boxes afterwards (collision.py:74-78)
        self._bounds_buf = _lib.alloc_bytes(ctx, max(self.n_nodes, 2) * 2 * 4 * self.program.coord_dtype.itemsize)
        # arrival words (upstream: one u32 per node); the fused build wants 2N words (N-1 arrival
        # words of the cross-window splits + N hand-over slots between its two passes)
        self._flags_buf = _lib.alloc_bytes(ctx, (self.n_nodes + 1) * 4)
        self._counts_buf = _lib.alloc_bytes(ctx, self.size * 4)
        self._scan_scratch = _lib.alloc_bytes(ctx, L.col_scan_scratch_bytes(self.size))
        self._total_buf = torch.zeros(1, dtype=torch.int64, device=ctx)
        self._wide_buf = _lib.alloc_bytes(ctx, max(self.size - 1, 1) * 32)   # 32-byte traversal records
        self._scene_buf = _lib.alloc_bytes(ctx, 8 * self.program.coord_dtype.itemsize)   # {min[4], max[4]} of the centres
        self._trv_scratch = None      # sized by the first call that writes rows (depends on its capacity)
        self._trv_scratch_bytes = 0
        self._bufs_for = key

    def resize(self, size=None, ngroups=None, group_size=None, radix_bits=None):
        if size is not None:
            self.size = size
        if group_size is not None:
            self.group_size = group_size
        # upstream passes roundUp(size, group_size) here (collision.py:94); the
        # sorter needs a multiple of 2 * group_size, which is what padded_size is
        self.sorter.resize(self.padded_size, group_size, radix_bits)
        self.reducer.resize(ngroups, group_size)

    @property
    def n_nodes(self):
        return self.size * 2 - 1

    @property
    def padded_size(self):
        return roundUp(self.size, 2 * self.group_size)

    # -- the hot path -----------------------------------------------------
    def get_collisions(self, cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf,
                       n_collisions, wait_for=None):
        if collisions_buf is None and n_collisions > 0:
            raise ValueError("Invalid collisions_buf for n_collisions > 0")
        self._ensure_scratch(cq)
        _lib.begin(cq, wait_for)
        if self.graph and self.stage_events is None and self._trv_marks is None:
            key = (cq.cuda_stream, self.size, n_collisions, _lib.ptr(coords_buf), _lib.ptr(radii_buf),
                   _lib.ptr(n_collisions_buf), _lib.ptr(collisions_buf))
            entry = self._graphs.get(key)
            if entry is None and self._graph_candidate == key:
                # second call with this signature: scratch is allocated, capture the sequence
                g = torch.cuda.CUDAGraph()
                before = self.kernel_launches
                with torch.cuda.graph(g, stream=cq):
                    self._enqueue(cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf, n_collisions)
                entry = self._graphs[key] = (g, self.kernel_launches - before)
                self.kernel_launches = before
                while len(self._graphs) > self.MAX_GRAPHS:     # a caller cycling through many buffers:
                    self._graphs.pop(next(iter(self._graphs)))  # forget the oldest capture
                # (buffers named in the key are referenced by the graph: the caller keeps them alive,
                #  as it must for any enqueued work)
            if entry is not None:
                with torch.cuda.stream(cq):
                    entry[0].replay()
                self.kernel_launches += entry[1]
                return _lib.done(cq)
            self._graph_candidate = key
        self._enqueue(cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf, n_collisions)
        return _lib.done(cq)

    def _enqueue(self, cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf, n_collisions):
        self._mark(cq, "begin", 0)
        self._sort_stage(cq, coords_buf)
        if self.fast:
            self._traverse_fast(cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf, n_collisions)
        else:
            self._traverse_reference_layout(cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf, n_collisions)

    def _mark(self, cq, name, kernels):
        self.kernel_launches += kernels
        if self.stage_events is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(cq)
            self.stage_events.append((name, ev))

    def _sort_stage(self, cq, coords_buf):
        """scene bounds -> Morton codes + ids -> stable sort (collision.py:157-169 upstream)."""
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr(cq)
        n, code, mark = self.size, self.program._code, self._mark
        # upstream keeps the scene bounds in slot 0 of the Bound array (dual use, collision.py:74-78);
        # the fast path needs them beyond the refit (grid of the traversal records), so they get a buffer
        scene = self._scene_buf if self.fast else self._bounds_buf
        _lib.check(L.col_bounds(P(coords_buf), n, code, 4, P(scene), P(self.reducer._ensure_scratch(cq)), S),
                   "col_bounds")
        mark(cq, "bounds", 1)
        _lib.check(L.col_morton(P(self._codes_bufs[0]), P(self._ids_bufs[0]), P(coords_buf), P(scene),
                                code, n, n, S), "col_morton")
        mark(cq, "morton", 1)
        self._sort_codes(cq)

    def _sort_codes(self, cq):
        """Stable sort of (_codes_bufs[0], _ids_bufs[0]) into (_codes_bufs[1], _ids_bufs[1])."""
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr(cq)
        n, mark = self.size, self._mark
        # padding keys are not needed by this sort: only the n real keys move
        _lib.check(L.col_sort_pairs(P(self._codes_bufs[0]), P(self._codes_bufs[1]), P(self._ids_bufs[0]),
                                    P(self._ids_bufs[1]), n, 4, 4, MORTON_BITS, P(self.sorter._ensure_scratch(cq=cq)), S),
                   "col_sort_pairs")
        mark(cq, "sort", 2 + (MORTON_BITS + 7) // 8)   # histogram + digit bases + one launch per digit

    def _build_wide(self, cq, coords_buf, radii_buf):
        """Karras tree + leaf boxes + refit, emitting the traversal records."""
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr(cq)
        n, code, mark = self.size, self.program._code, self._mark
        if self.fused_build:
            _lib.check(L.col_build_refit_wide(P(self._codes_bufs[1]), P(self._ids_bufs[1]), P(self._nodes_buf),
                                              P(self._wide_buf), P(self._flags_buf), P(self._bounds_buf),
                                              P(coords_buf), P(radii_buf), P(self._scene_buf), code, n, S),
                       "col_build_refit_wide")
            mark(cq, "tree+refit", 2 if n >= (1 << 22) else 1)   # (large scenes: window pass + top pass)
            return
        _lib.check(L.col_build_tree_wide(P(self._codes_bufs[1]), P(self._ids_bufs[1]), P(self._nodes_buf),
                                         P(self._wide_buf), P(self._flags_buf), n, S), "col_build_tree_wide")
        mark(cq, "tree", 1)
        _lib.check(L.col_refit_wide(P(self._bounds_buf), P(self._wide_buf), P(self._flags_buf), P(self._nodes_buf),
                                    P(coords_buf), P(radii_buf), P(self._scene_buf), code, n, S), "col_refit_wide")
        mark(cq, "refit", 1)

    def _walk_range(self, cq, collisions_buf, n_collisions, q_begin, q_end, base=0):
        """Queries [q_begin, q_end) (sorted positions): every query walks once (hits
        parked), the counts are scanned, the rows are emitted in the deterministic
        order.  Leaves the u64 row count of the range in self._total_buf."""
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr(cq)
        need = L.col_traverse_scratch_bytes(q_end - q_begin, n_collisions)
        if need > self._trv_scratch_bytes:   # (grows with the rows per query: deeper per-query stash)
            with torch.cuda.stream(cq):      # allocated in cq's order, like the rest of the scratch
                self._trv_scratch = _lib.alloc_bytes(self.program.context, need)
            self._trv_scratch_bytes = need
            self._graphs.clear()
            self._graph_candidate = None
        args = (P(collisions_buf), n_collisions, base, P(self._total_buf), P(self._wide_buf), P(self._bounds_buf),
                P(self._scene_buf), P(self._ids_bufs[1]), self.program._code, self.size, q_begin, q_end,
                P(self._trv_scratch), S)
        if self._trv_marks is not None and n_collisions:
            _lib.check(L.col_traverse_wide_pairs_marked(*args, self._trv_marks), "col_traverse_wide_pairs_marked")
        else:
            _lib.check(L.col_traverse_wide_pairs(*args), "col_traverse_wide_pairs")
        self._mark(cq, "traverse_pairs" if n_collisions else "traverse_count", 4 if n_collisions else 1)

    def _traverse_fast(self, cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf, n_collisions):
        self._build_wide(cq, coords_buf, radii_buf)
        self._walk_range(cq, collisions_buf, n_collisions, 0, self.size)
        # counter = true total (low 32 bits, as upstream's uint counter)
        _lib.check(_lib.lib().col_copy(_lib.ptr(n_collisions_buf), _lib.ptr(self._total_buf), 4,
                                       _lib.stream_ptr(cq)), "col_copy")

    def _traverse_reference_layout(self, cq, coords_buf, radii_buf, n_collisions_buf, collisions_buf, n_collisions):
        """The same stages on the reference's Node / Bound arrays only (kernel-level
        contract of collision.cl); count -> scan -> write, or one atomic counter."""
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr(cq)
        n, code, mark = self.size, self.program._code, self._mark
        _lib.check(L.col_build_tree(P(self._codes_bufs[1]), P(self._ids_bufs[1]), P(self._nodes_buf), n, S),
                   "col_build_tree")
        mark(cq, "tree", 1)
        _lib.check(L.col_refit(P(self._bounds_buf), P(self._flags_buf), P(self._nodes_buf), P(coords_buf),
                               P(radii_buf), code, n, S), "col_refit")
        mark(cq, "refit", 1)
        if self.ordered:
            _lib.check(L.col_traverse_count(P(self._counts_buf), P(self._nodes_buf), P(self._bounds_buf), code,
                                            n, 0, n, S), "col_traverse_count")
            mark(cq, "traverse_count", 1)
            _lib.check(L.col_exclusive_scan_u32(P(self._counts_buf), n, P(self._total_buf),
                                                P(self._scan_scratch), S), "col_exclusive_scan_u32")
            # counter = true total (low 32 bits, as upstream's uint counter)
            _lib.check(L.col_copy(P(n_collisions_buf), P(self._total_buf), 4, S), "col_copy")
            mark(cq, "scan", 1)
            if n_collisions > 0:
                _lib.check(L.col_traverse_write(P(collisions_buf), n_collisions, P(self._counts_buf), 0,
                                                P(self._nodes_buf), P(self._bounds_buf), code, n, 0, n, S),
                           "col_traverse_write")
                mark(cq, "traverse_write", 1)
        else:
            _lib.check(L.col_fill_u32(P(n_collisions_buf), 0, 1, S), "col_fill_u32")
            _lib.check(L.col_traverse_atomic(P(collisions_buf), P(n_collisions_buf), n_collisions,
                                             P(self._nodes_buf), P(self._bounds_buf), code, n, 0, n, S),
                       "col_traverse_atomic")
            mark(cq, "traverse_atomic", 1)
