/*
 * collision_b200.h -- C ABI of libcollision_b200.so
 *
 * B200 (sm_100a) implementation of the LBVH sphere-collision hot path of
 * kwohlfahrt/collision.  The reference has no FFI of its own (its boundary is
 * Python over PyOpenCL), so every entry point below cites the reference kernel
 * / enqueue site it replaces (paths relative to the reference checkout).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no C++/torch types.
 *   - Every pointer is a DEVICE pointer unless its name starts with h_.
 *   - Every call only ENQUEUES work on `stream` (a cudaStream_t passed as
 *     void*): it never allocates, never synchronises, never reads back.
 *   - Scratch is caller-provided; its size comes from the pure host function
 *     *_scratch_bytes.  Scratch needs no initialisation by the caller.
 *   - Return value: 0 on success, a cudaError_t value on a CUDA failure,
 *     COL_ERR_* (negative) on a bad argument.  Nothing throws.
 *   - Re-entrant for distinct (stream, scratch) pairs; no global state.
 *   - Floating-point work is compiled with -fmad=false -prec-div=true
 *     -ftz=false: IEEE, no contraction, correctly rounded division.
 */
#ifndef COLLISION_B200_H
#define COLLISION_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define COL_F32 0
#define COL_F64 1
/* integer value types of col_sum (tests/test_summer_py.py:12 sums uint32x4 and int32) */
#define COL_U32 2
#define COL_I32 3
#define COL_U64 4
#define COL_I64 5

#define COL_ERR_BAD_ARG (-1)
#define COL_ERR_TOO_LARGE (-2)
#define COL_ERR_UNSUPPORTED (-3)

/* ABI version of this header; bumped on any signature change. */
int col_abi_version(void);
/* Compute capability the kernels were built for (100 -> sm_100a). */
int col_built_arch(void);

/* ---------------------------------------------------------------- layouts
 * struct Node (collision/collision.cl:42-53; numpy dtype collision/collision.py:9):
 *   16 bytes {u32 parent; u32 right_edge; u32 data[2]} -- leaf: data[0] = id,
 *   internal: data = children[2].  Array of 2n-1: internals [0, n-1), leaves
 *   [n-1, 2n-1), node 0 = root.  NO_NODE = 0xFFFFFFFF.
 * struct Bound (collision/collision.cl:123-126): {T min[4]; T max[4]} (3-vector
 *   in 4 lanes): 32 bytes for f32, 64 bytes for f64; lane 3 is never meaningful.
 * Coordinates: n slots of 4 lanes (xyz + ignored pad), collision/misc.py:62-68.
 */
typedef struct {
    uint32_t parent;
    uint32_t right_edge;
    uint32_t data[2];
} col_node;

/* ----------------------------------------------------------------- reduce
 * Replaces Reducer.reduce = bounds1 + bounds2 (collision/reduce.py:62-76,
 * collision/reduce.cl:5-58) with the accumulator list of Bounds
 * (collision/bounds.py:5): out_minmax = {min[lanes], max[lanes]} over n slots of
 * `lanes` scalars (1, 2, 4, 8 or 16; a 3-vector is a 4-lane slot).  One launch.
 * col_sum is the Summer accumulator (collision/summer.py:4-8): out = {sum[lanes]};
 * besides COL_F32/COL_F64 it takes the integer codes (sums wrap, exact in any order). */
size_t col_reduce_scratch_bytes(void);
int col_bounds(const void* values, uint64_t n, int dtype, uint32_t lanes, void* out_minmax,
               void* scratch, void* stream);
int col_sum(const void* values, uint64_t n, int dtype, uint32_t lanes, void* out_sum,
            void* scratch, void* stream);

/* ----------------------------------------------------------------- Morton
 * Replaces range (collision/collision.cl:8-10), the 0xFFFFFFFF padding fill
 * (collision/collision.py:138-142) and calculateCodes/morton/expandBits
 * (collision/collision.cl:14-40) in one pass: codes[i] for i < n, codes[i] =
 * 0xFFFFFFFF for n <= i < n_padded, ids[i] = i for i < n_padded.  ids may be
 * NULL (codes only, the kernel-level calculateCodes contract).
 * range2x4 = {min[4], max[4]} as written by col_bounds. */
int col_morton(uint32_t* codes, uint32_t* ids, const void* coords4, const void* range2x4,
               int dtype, uint32_t n, uint32_t n_padded, void* stream);

/* ------------------------------------------------------------------- scan
 * Replaces PrefixScanner.prefix_sum (collision/scan.py:75-112; local_scan /
 * block_scan, collision/scan.cl:5-36): in-place exclusive prefix sum of u32,
 * wrapping mod 2^32, single pass (decoupled look-back).  If total_out is not
 * NULL it receives the u64 (non-wrapping) grand total. */
size_t col_scan_scratch_bytes(uint64_t n);
int col_exclusive_scan_u32(uint32_t* data, uint64_t n, uint64_t* total_out, void* scratch,
                           void* stream);
/* The same scan with NO caller scratch and no allocation of any kind (no cudaMalloc / cudaFree,
 * safe under CUDA graph capture and in loops that must not allocate).  This entry point has no
 * counterpart in the reference.  The ticket
 * and status words live in a 16,640-byte device array built into the library (every GPU that loads
 * the library holds it until exit), which limits n to COL_SCAN_STATIC_MAX_N values: larger n is
 * refused with COL_ERR_TOO_LARGE and a message on stderr.  All calls on one GPU share the array:
 * calls that overlap on the GPU (different streams or threads) race -- serialise them.  Results
 * are identical to col_exclusive_scan_u32 (integer arithmetic, bit-exact). */
#define COL_SCAN_STATIC_MAX_N (1ull << 24)
int col_exclusive_scan_u32_static(uint32_t* data, uint64_t n, uint64_t* total_out, void* stream);

/* ------------------------------------------------------------------- sort
 * Replaces RadixSorter.sort (collision/radix.py:118-170; block_sort / scatter,
 * collision/radix.cl:48-139) and its per-pass scans: STABLE LSD radix sort
 * (8-bit digits, one-sweep with decoupled look-back) of n keys of key_bytes
 * (4 or 8) with an optional payload of value_bytes (0, 4, 8, 16 or 32) per key.
 * keys_in / vals_in are NOT modified (the reference clobbers them).  Only key
 * bits [0, end_bit) are sorted on (end_bit = 0 -> all key bits); the Collider
 * passes 30 because Morton codes are 30-bit.  n <= 2^30. */
size_t col_sort_scratch_bytes(uint64_t n, uint32_t key_bytes, uint32_t value_bytes);
int col_sort_pairs(const void* keys_in, void* keys_out, const void* vals_in, void* vals_out,
                   uint64_t n, uint32_t key_bytes, uint32_t value_bytes, uint32_t end_bit,
                   void* scratch, void* stream);
/* One LSD pass in isolation (what block_sort + prefix_sum + scatter compute per
 * pass, collision/radix.py:140-157): stable partition by the digit
 * (key >> (pass * radix_bits)) & (2^radix_bits - 1), radix_bits <= 8. */
int col_radix_pass(const void* keys_in, void* keys_out, const void* vals_in, void* vals_out,
                   uint64_t n, uint32_t key_bytes, uint32_t value_bytes, uint32_t radix_bits,
                   uint32_t pass, void* scratch, void* stream);

/* ------------------------------------------------------------------- tree
 * Replaces fillInternal + generateBVH/delta (collision/collision.cl:55-121) in
 * one launch: writes all 2n-1 col_node records (leaves: id, right_edge;
 * internals: children, right_edge; every non-root parent).  The two fields the
 * reference never writes are set deterministically: nodes[0].parent = NO_NODE,
 * leaf data[1] = NO_NODE.  ids_sorted may be NULL (internals + parents only,
 * the kernel-level generateBVH contract; leaf id/right_edge left untouched). */
int col_build_tree(const uint32_t* codes_sorted, const uint32_t* ids_sorted, void* nodes,
                   uint32_t n, void* stream);
/* fillInternal alone (collision/collision.cl:55-63). */
int col_fill_leaves(void* nodes, const uint32_t* ids_sorted, uint32_t n, void* stream);

/* ------------------------------------------------------------------ refit
 * Replaces the flag clear (collision/collision.py:147-150), leafBounds and
 * internalBounds (collision/collision.cl:128-162) in one launch after a flag
 * memset: leaf boxes = coords[id] -/+ radii[id]; every leaf walks to the root,
 * the second arrival at a node (atomic flag) merges the two child boxes.
 * flags: 2n-1 u32 of scratch (cleared inside).  coords4/radii may be NULL to
 * skip the leaf-box step (kernel-level internalBounds contract). */
int col_refit(void* bounds, uint32_t* flags, const void* nodes, const void* coords4,
              const void* radii, int dtype, uint32_t n, void* stream);
int col_leaf_bounds(void* bounds, const void* coords4, const void* radii, const void* nodes,
                    int dtype, uint32_t n, void* stream);

/* --------------------------------------------------------------- traverse
 * Replaces traverse/checkOverlap/isLeaf (collision/collision.cl:164-226).
 *
 * col_traverse_count: one thread per query leaf q in [q_begin, q_end) (sorted
 * position): counts[q - q_begin] = number of reported pairs of that query.
 *
 * col_traverse_write: the same walk, writing row (id[q], id[leaf]) at
 * offsets[q - q_begin] + k for the k-th hit of q in DFS order (child a before
 * b) when that index < capacity: the deterministic order of the sequential
 * work-item loop.  offsets = exclusive scan of counts (+ a base of the caller's
 * choice, e.g. the pairs of lower ranks).
 *
 * col_traverse_atomic: the reference's own single-pass contract: rows appended
 * through one global counter in ARBITRARY order, *counter ends as the true
 * total even beyond capacity (collision.cl:202-204).  pairs may be NULL when
 * capacity == 0 (count only).  counter must be zeroed by the caller.  */
int col_traverse_count(uint32_t* counts, const void* nodes, const void* bounds, int dtype,
                       uint32_t n, uint32_t q_begin, uint32_t q_end, void* stream);
int col_traverse_write(uint32_t* pairs, uint64_t capacity, const uint32_t* offsets,
                       uint64_t base, const void* nodes, const void* bounds, int dtype,
                       uint32_t n, uint32_t q_begin, uint32_t q_end, void* stream);
int col_traverse_atomic(uint32_t* pairs, uint32_t* counter, uint32_t capacity, const void* nodes,
                        const void* bounds, int dtype, uint32_t n, uint32_t q_begin,
                        uint32_t q_end, void* stream);

/* ------------------------------------------------- Collider fast path ("wide")
 * The entry points above keep the reference's Node/Bound kernel contract.  The
 * Collider itself (collision/collision.py:171-196: fillInternal, generateBVH,
 * leafBounds, internalBounds, traverse) runs on a traversal-oriented copy of the
 * tree that the build and refit kernels emit alongside the Node/Bound arrays:
 * one 32-byte record per internal node (any coordinate dtype):
 *   u32 child a, u32 child b, 6 x u16 box of child a, 6 x u16 box of child b,
 *   each box as {lo.x lo.y | lo.z ~hi.x | ~hi.y ~hi.z} (~v = 65535 - v) so that the box
 *   test is three packed "record <= query" comparisons (max.u16x2)
 * The u16 boxes live on a 65536^3 grid over the scene bounds and are rounded
 * OUTWARD, so an internal node is never pruned wrongly (a few extra nodes may be
 * visited); whether a LEAF is reported is always decided on the exact float boxes
 * of the Bound array, exactly as collision.cl:164-166 -- the reported rows and
 * their order are identical to the entry points above.  A visit is one 256-bit load.
 *
 * `wide`: 32 * (n-1) bytes.  `flags`: n-1 u32 arrival words, initialised by
 * col_build_tree_wide and consumed by col_refit_wide (call them as a pair).
 * `scene2x4`: {min[4], max[4]} of the centres as written by col_bounds -- NOT slot
 * 0 of `bounds` (the refit overwrites that with the root box).
 *
 * col_traverse_wide_pairs: one walk per query + deterministic compaction (walk
 * and park hits -> single-pass scan of the counts -> emit): row k of query q
 * lands at base + (number of rows of queries q_begin..q-1) + k when that index
 * is < capacity; *total_out (u64, required) = rows found in the range (the list
 * positions are exact for totals < 2^32 per call).  capacity == 0: count only.
 * scratch: col_traverse_scratch_bytes(q_end - q_begin, capacity) -- the per-query stash is
 * 32, 64 or 128 rows deep depending on the rows per query the caller made room for; a
 * scratch sized for a capacity serves every call with that capacity or less.
 * ids_sorted: ids in sorted order (the sort's payload output). */
int col_build_tree_wide(const uint32_t* codes_sorted, const uint32_t* ids_sorted, void* nodes, void* wide,
                        uint32_t* flags, uint32_t n, void* stream);
int col_refit_wide(void* bounds, void* wide, uint32_t* flags, const void* nodes, const void* coords4,
                   const void* radii, const void* scene2x4, int dtype, uint32_t n, void* stream);
/* col_build_tree_wide + col_refit_wide as ONE bottom-up climb (refit.cu): the radix tree is found
 * from the leaves upwards in the same climb that merges the boxes -- a pass over the 256-leaf
 * windows (shared-memory arrivals), then a pass over the few nodes that span windows.  Node
 * array, Bound array and traversal records are bit-identical to the pair above.
 * `flags`: 2n u32 of scratch, set by the call (n-1 arrival words + n hand-over slots). */
int col_build_refit_wide(const uint32_t* codes_sorted, const uint32_t* ids_sorted, void* nodes, void* wide,
                         uint32_t* flags, void* bounds, const void* coords4, const void* radii,
                         const void* scene2x4, int dtype, uint32_t n, void* stream);
size_t col_traverse_scratch_bytes(uint32_t n_queries, uint64_t capacity);
int col_traverse_wide_count(uint64_t* total, uint32_t* counts, const void* wide, const void* bounds,
                            const void* scene2x4, int dtype, uint32_t n, uint32_t q_begin, uint32_t q_end,
                            void* stream);
int col_traverse_wide_pairs(uint32_t* pairs, uint64_t capacity, uint64_t base, uint64_t* total_out,
                            const void* wide, const void* bounds, const void* scene2x4,
                            const uint32_t* ids_sorted, int dtype, uint32_t n, uint32_t q_begin,
                            uint32_t q_end, void* scratch, void* stream);
/* The same call with optional stage marks: h_marks3 is a HOST array of three cudaEvent_t handles
 * (any may be NULL) recorded on `stream` after the walk, after the exact confirmation and after
 * the scan of the row counts -- what follows the third is the emit.  Measurement aid (bench.py's
 * "compaction" figure: scan + emit); it launches exactly the kernels of col_traverse_wide_pairs. */
int col_traverse_wide_pairs_marked(uint32_t* pairs, uint64_t capacity, uint64_t base, uint64_t* total_out,
                                   const void* wide, const void* bounds, const void* scene2x4,
                                   const uint32_t* ids_sorted, int dtype, uint32_t n, uint32_t q_begin,
                                   uint32_t q_end, void* scratch, void* stream, void* const* h_marks3);
int col_traverse_wide_write(uint32_t* pairs, uint64_t capacity, const uint32_t* offsets, uint64_t base,
                            const void* wide, const void* bounds, const void* scene2x4,
                            const uint32_t* ids_sorted, int dtype, uint32_t n, uint32_t q_begin,
                            uint32_t q_end, void* stream);

/* ----------------------------------------------------------- find_offsets
 * Replaces OffsetFinder.find_offsets (collision/offset.py:40-49,
 * collision/offset.cl:3-12): for sorted unsigned values[n_values],
 * offsets[v] = first i with values[i] >= v, for v in [0, n_offsets); entries
 * above the maximum = n_values.  value_bytes / offset_bytes: 4 or 8. */
int col_find_offsets(const void* values, uint64_t n_values, uint32_t value_bytes, void* offsets,
                     uint64_t n_offsets, uint32_t offset_bytes, void* stream);

/* --------------------------------------------------------- gather/scatter
 * Replaces Indexer.gather / .scatter (collision/index.py:37-55,
 * collision/index.cl:1-13): out[i] = in[idx[i]] / out[idx[i]] = in[i];
 * value_bytes any multiple of 4 up to 128 (or 1, 2), index_bytes 4 or 8. */
int col_gather(const void* in, const void* idx, void* out, uint64_t n, uint32_t value_bytes,
               uint32_t index_bytes, void* stream);
int col_scatter(const void* in, const void* idx, void* out, uint64_t n, uint32_t value_bytes,
                uint32_t index_bytes, void* stream);

/* ------------------------------------------- kernel-level contracts (compat.cu)
 * Upstream's tests launch the kernels of the non-collision programs directly, with the
 * reference's own launch geometry ("the kernel-level argument/layout
 * contract is public too").  These entry points stand behind `program.kernels[name]` of
 * RadixProgram, PrefixScanProgram, ReductionProgram and OffsetProgram; the hot path
 * (col_sort_pairs, col_exclusive_scan_u32, col_bounds, col_find_offsets) never calls them.
 * One CUDA block per OpenCL work-group; `items_per_group` = 2 * local_size.
 *
 * col_compat_local_scan   local_scan   collision/scan.cl:5-30: every group's exclusive scan in
 *                         place, group totals to block_sums[group] (block_sums may be NULL).
 * col_compat_block_scan   block_scan   collision/scan.cl:32-36: data[i] += block_sums[group of i].
 * col_compat_block_sort   block_sort   collision/radix.cl:48-102: every group stable-sorted by
 *                         the digit in place (values, when given, move with their keys; value_bytes
 *                         a multiple of 4); histogram[digit * n_groups + group] = the group's counts.
 * col_compat_scatter      scatter      collision/radix.cl:104-139: block-sorted groups to
 *                         out[offset[d * n_groups + g] + i - first slot of digit d in the group].
 * col_compat_find_offsets find_offsets collision/offset.cl:3-12, literally (thread g: offsets[i] = g + 1
 *                         for i in (values[g], values[g + 1]]; offsets[g] = 0 for g <= values[0]); the
 *                         fill is the host's (collision/offset.py:43-46).
 * col_compat_reduce1/2    bounds1 / bounds2  collision/reduce.cl:5-58 for ANY accumulator list
 *                         (collision/reduce.py:9-22): h_ops[a] in {0: min, 1: max, 2: ADD}, h_inits[a] the
 *                         initialiser (HOST arrays, n_acc <= 8); reduce1 = ngroups x group_size
 *                         threads, grid-stride, group_accs[group][acc][lane]; reduce2 = one group of
 *                         `ngroups` threads folding the partials IN PLACE with upstream's halving loop
 *                         -- a non-power-of-two ngroups drops the tail partials exactly as upstream
 *                         (reduce.cl:42-52).  dtype: COL_F32 ... COL_I64; lanes <= 16. */
int col_compat_local_scan(uint32_t* data, uint32_t* block_sums, uint64_t n_groups, uint32_t items_per_group,
                          void* stream);
int col_compat_block_scan(uint32_t* data, const uint32_t* block_sums, uint64_t n_items, uint32_t items_per_group,
                          void* stream);
int col_compat_block_sort(void* keys, void* values, uint32_t* histogram, uint64_t n_groups, uint32_t items_per_group,
                          uint32_t key_bytes, uint32_t value_bytes, uint32_t radix_bits, uint32_t pass, void* stream);
int col_compat_scatter(const void* keys, void* out_keys, const void* values, void* out_values, const uint32_t* offset,
                       const uint32_t* histogram, uint64_t n_groups, uint32_t items_per_group, uint32_t key_bytes,
                       uint32_t value_bytes, uint32_t radix_bits, uint32_t pass, void* stream);
int col_compat_find_offsets(const void* values, uint64_t n_threads, uint32_t value_bytes, void* offsets,
                            uint32_t offset_bytes, void* stream);
int col_compat_reduce1(const void* values, uint64_t n, int dtype, uint32_t lanes, const int* h_ops,
                       const double* h_inits, uint32_t n_acc, void* group_accs, uint32_t ngroups, uint32_t group_size,
                       void* stream);
int col_compat_reduce2(void* group_accs, void* output, int dtype, uint32_t lanes, const int* h_ops, uint32_t n_acc,
                       uint32_t ngroups, void* stream);

/* ------------------------------------------------ buffers shared across ranks (ipc.cu)
 * Multi-GPU only (no reference counterpart: upstream is single-device).  One process per GPU;
 * a rank exports a cudaMalloc allocation as a 64-byte CUDA IPC handle (host memory), the
 * other ranks of the same box open it on their own device and read it from their kernels
 * over NVLink.  These are the only entry points that allocate. */
int col_ipc_alloc(void** ptr, uint64_t bytes);
int col_ipc_free(void* ptr);
int col_ipc_export(void* ptr, void* h_handle64);
int col_ipc_open(void** ptr, const void* h_handle64);
int col_ipc_close(void* ptr);

/* ------------------------------------------------ sharded build (multi-GPU, refit.cu)
 * No reference counterpart.  The windows (256 sorted leaves each) are dealt out to the ranks:
 * col_shard_build_windows  = col_build_refit_wide's window pass for windows [win_begin, win_end) only; the
 *                            subtree roots it hands over are packed into `packed` ({count, overflow} header +
 *                            up to `cap` records {node, lo, hi, codes at the ends, box}; col_shard_packed_bytes(cap,
 *                            dtype) bytes): the rank's PROBE list (it stays on the rank).
 * col_shard_top_local      = the nodes above the windows that still lie inside the rank's own range of leaves
 *                            [leaf_begin, leaf_end): climbs from the handed-over roots while the parent stays in
 *                            range.  Fills `exchange` (col_shard_exchange_bytes(span_cap, dtype) bytes): a header
 *                            {roots, overflow, spanning nodes, overflow}, the roots of the rank's maximal in-range
 *                            subtrees (a few dozen) and the indices of the up to span_cap nodes it finished.
 *                            The caller all-gathers the ranks' exchange buffers (rank-major) into `gathered`.
 * col_shard_top            = the handful of nodes that span RANKS, from the gathered range roots (on every rank),
 *                            and the other ranks' spanning records copied from their HBM by the gathered lists
 *                            (wide_table: the ranks' record arrays as mapped into this process, col_ipc_open).
 *                            The codes at the other ranks' roots' ends are written into codes_sorted.
 * col_shard_halo           = which in-window subtrees of OTHER ranks can this rank's queries enter (a walk of
 *                            the spanning nodes with each root's box of the probe list), and their records + leaf
 *                            boxes copied from the owners' buffers (wide_table / bounds_table) into `wide` / `bounds`.
 *                            win_edges: world + 1 window indices (device), rank r owns [win_edges[r], win_edges[r+1]);
 *                            list: col_shard_list_bytes(list_cap) bytes of scratch.  After it the traversal of
 *                            the rank's own queries [leaf_begin, leaf_end) finds everything it touches locally.
 *                            ids_table / ids_sorted (may both be NULL): the sphere ids of the fetched leaves are
 *                            copied from the owners as well (needed after col_shard_sort, see below).
 * Overflow of `cap` / span_cap / `list_cap` sets word 1 / 3 of the respective header (the host checks them when it
 * next syncs); when any rank's exchange overflowed, col_shard_top writes a root that ends every walk at once and the
 * halo asks for nothing, so that nothing ever walks half-built nodes.
 *
 * col_shard_sort           = the SORT dealt out the same way: positions [pos_lo, pos_hi] of the stable sort of all
 *                            n codes (ids = input positions), written at codes_sorted / ids_sorted + those positions
 *                            and nowhere else -- bit-identical to what col_sort_pairs writes there.  The codes' top
 *                            8-bit digit is counted, the buckets covering the positions are selected ON THE DEVICE
 *                            (range_out: 4 words {offset, count, first bucket, last bucket}; no host round trip),
 *                            their (code, index) pairs are collected in input order and sorted on end_bit bits.
 *                            `codes` must be 16-byte aligned; scratch: col_shard_sort_scratch_bytes(n).
 *                            The records a rank hands over carry the codes at their ends, and col_shard_top writes
 *                            those of the other ranks into codes_sorted (all the top pass reads), so a rank never
 *                            needs the rest of the sorted order. */
size_t col_shard_packed_bytes(uint32_t cap, int dtype);
size_t col_shard_list_bytes(uint32_t list_cap);
int col_shard_build_windows(const uint32_t* codes_sorted, const uint32_t* ids_sorted, void* nodes, void* wide,
                            uint32_t* flags, void* bounds, const void* coords4, const void* radii,
                            const void* scene2x4, int dtype, uint32_t n, uint32_t win_begin, uint32_t win_end,
                            void* packed, uint32_t cap, void* stream);
size_t col_shard_exchange_bytes(uint32_t span_cap, int dtype);
int col_shard_top_local(const uint32_t* codes_sorted, void* nodes, void* wide, uint32_t* flags, void* bounds,
                        const void* scene2x4, int dtype, uint32_t n, uint32_t leaf_begin, uint32_t leaf_end,
                        const void* probe, uint32_t probe_cap, void* exchange, uint32_t span_cap, void* stream);
int col_shard_top(const void* gathered, uint32_t span_cap, uint32_t world, uint32_t my_rank,
                  uint32_t* codes_sorted, void* nodes, void* wide, uint32_t* flags, void* bounds,
                  const void* scene2x4, int dtype, uint32_t n, const uint64_t* wide_table, void* stream);
int col_shard_halo(const void* probe, uint32_t cap, const void* gathered, uint32_t span_cap, uint32_t world,
                   uint32_t my_rank, void* wide, void* bounds,
                   const void* scene2x4, int dtype, uint32_t* flags, void* list, uint32_t list_cap,
                   const uint64_t* wide_table, const uint64_t* bounds_table, const uint32_t* win_edges,
                   uint32_t n, uint32_t leaf_begin, uint32_t leaf_end, const uint64_t* ids_table,
                   uint32_t* ids_sorted, void* stream);
/* The same selection with the data EXCHANGED instead of replicated (no rank ever holds all the codes or the whole
 * scene): col_shard_count = top-digit counts of the rank's own slice of the codes (hist256; the caller all-gathers
 * them, rank-major, into hist_all); col_shard_push = the plan every rank derives from hist_all (plan: 8 words per
 * rank {offset, count, first bucket, last bucket, where this rank's entries start in that rank's receive buffers})
 * and, per destination, a stable compaction of the own slice's {code, sphere id = id_base + i, coordinates, radius}
 * written straight into the destination's receive buffers (h_recv_*: HOST arrays of `world` device addresses as
 * mapped into this process, col_ipc_open; the rank's own entry = its own buffers; each sized for n entries);
 * pos_table: world x {pos_lo, pos_hi} (device).  The caller then runs any collective on the stream (all pushes have
 * landed once every rank has passed it) and calls col_shard_sort_received: stable sort of the received codes --
 * clobbered -- placed at codes_sorted + offset, with place_sorted = each entry's index in the receive buffers (the
 * build reads coordinates and radii through it) and ids_sorted = the sphere ids; range = the rank's own plan entry.
 * Positions [offset, offset + count) come out bit-identical to col_sort_pairs over the whole scene. */
int col_shard_count(const uint32_t* codes_own, uint32_t n_own, uint32_t end_bit, uint32_t* hist256, void* stream);
size_t col_shard_push_scratch_bytes(uint64_t n_own, uint32_t world);
int col_shard_push(const uint32_t* hist_all, uint32_t world, uint32_t my_rank, const uint32_t* pos_table,
                   uint32_t n, const uint32_t* codes_own, uint32_t n_own, uint32_t id_base,
                   const void* coords_own, const void* radii_own, int dtype, uint32_t end_bit,
                   const uint64_t* h_recv_codes, const uint64_t* h_recv_ids, const uint64_t* h_recv_coords,
                   const uint64_t* h_recv_radii, uint32_t* plan, void* scratch, void* stream);
size_t col_shard_sort_received_scratch_bytes(uint64_t n);
int col_shard_sort_received(uint32_t* recv_codes, const uint32_t* recv_ids, uint32_t n, const uint32_t* range,
                            uint32_t end_bit, uint32_t* codes_sorted, uint32_t* place_sorted, uint32_t* ids_sorted,
                            void* scratch, void* stream);
size_t col_shard_sort_scratch_bytes(uint64_t n);
int col_shard_sort(const uint32_t* codes, uint32_t n, uint32_t pos_lo, uint32_t pos_hi, uint32_t end_bit,
                   uint32_t* codes_sorted, uint32_t* ids_sorted, uint32_t* range_out, void* scratch,
                   void* stream);

/* utility: u32 fill (cl.enqueue_fill_buffer sites, e.g. collision/offset.py:43) */
int col_fill_u32(uint32_t* dst, uint32_t value, uint64_t n, void* stream);
/* range (collision/collision.cl:8-10): ids[i] = i for i < n */
int col_range(uint32_t* ids, uint64_t n, void* stream);
/* device-to-device copy of `bytes` (cl.enqueue_copy sites) */
int col_copy(void* dst, const void* src, uint64_t bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* COLLISION_B200_H */
