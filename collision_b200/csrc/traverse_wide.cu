// traverse_wide.cu -- the Collider's traversal: wide nodes + single-pass
// deterministic pair compaction.
//
// Same walk as collision/collision.cl:174-226 (see traverse.cu, which keeps the
// reference's Node/Bound argument contract), restructured around HBM/L1 traffic:
//
//  * one 32-byte record per internal node (qnode.cuh) holds both children's indices and
//    their boxes on a conservative u16 grid: a visit is ONE 256-bit load instead of seven
//    128-bit loads from five lines (ncu: the walk is bound by L1 data-pipe wavefronts, one
//    per diverged lane per load).  Leaves are confirmed on the exact float boxes, so the
//    reported rows and their order are the reference's, bit for bit.
//  * count -> scan -> write without a second walk: every query walks once and
//    parks its first `depth` hits (the stash: 32..128 rows) in a k-major global array (coalesced rows, only
//    the used ones move); the counts are scanned by the single-pass scan kernel;
//    an emit kernel copies the parked hits to their final rows.  The row order is
//    exactly that of the reference's work-items run sequentially (query-major,
//    DFS order, child a before b).  Only queries with more than `depth` hits walk
//    a second time, for the remainder.  (A single fused kernel chaining blocks by
//    decoupled look-back was measured slower: blocks of dense regions stall every
//    later block, and a shared-memory stash shrinks the L1 the walk depends on.)

#include "common.cuh"
#include "qnode.cuh"

namespace {

#ifndef COL_TRV_THREADS
#define COL_TRV_THREADS 256
#endif
constexpr int TRV_THREADS = COL_TRV_THREADS;
constexpr uint32_t TRV_STASH_MIN = 32, TRV_STASH_MAX = 128;
constexpr uint32_t TRV_DONE = 0xFFFFFFFFu;   // stack sentinel / "lane has no node to visit"

struct QNode { uint32_t w[8]; };   // see qnode.cuh

__device__ __forceinline__ QNode load_qnode(const unsigned char* __restrict__ wide, uint32_t idx) {
    QNode q;
#ifndef COL_QNODE_LD128   // (two 128-bit loads measured 3 % slower on B200)
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(q.w[0]), "=r"(q.w[1]), "=r"(q.w[2]), "=r"(q.w[3]), "=r"(q.w[4]), "=r"(q.w[5]), "=r"(q.w[6]),
                   "=r"(q.w[7])
                 : "l"(wide + (size_t)idx * QNODE_BYTES));
#else
    const uint4* p = reinterpret_cast<const uint4*>(wide + (size_t)idx * QNODE_BYTES);
    const uint4 a = __ldg(p), b = __ldg(p + 1);
    q.w[0] = a.x; q.w[1] = a.y; q.w[2] = a.z; q.w[3] = a.w;
    q.w[4] = b.x; q.w[5] = b.y; q.w[6] = b.z; q.w[7] = b.w;
#endif
    return q;
}

// exact query box + its outward-rounded grid image (query side of the packed test)
template <typename T> struct Query {
    T lo[3], hi[3];
    uint3 grid;
};

template <typename T> struct LeafBox { T lo[3], hi[3]; };

__device__ __forceinline__ LeafBox<float> load_leaf_box(const float* __restrict__ bounds, uint32_t node) {
    float v[8];
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(bounds + (size_t)node * 8));
    return {{v[0], v[1], v[2]}, {v[4], v[5], v[6]}};
}
__device__ __forceinline__ LeafBox<double> load_leaf_box(const double* __restrict__ bounds, uint32_t node) {
    const double2* p = reinterpret_cast<const double2*>(bounds) + (size_t)node * 4;
    const double2 a = __ldg(p), c = __ldg(p + 1), d = __ldg(p + 2), e = __ldg(p + 3);
    return {{a.x, a.y, c.x}, {d.x, d.y, e.x}};
}

template <typename T>
__device__ __forceinline__ Query<T> load_query(const T* __restrict__ bounds, uint32_t node, const Quant<T>& qz) {
    const LeafBox<T> b = load_leaf_box(bounds, node);
    Query<T> q;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        q.lo[k] = b.lo[k];
        q.hi[k] = b.hi[k];
    }
    q.grid = quantise_query(qz, q.lo, q.hi);
    return q;
}

// collision.cl:164-166: all(q.max > c.min & q.min < c.max), strict, x y z -- on the exact boxes
template <typename T>
__device__ __forceinline__ bool exact_overlap(const Query<T>& q, const LeafBox<T>& c) {
    return (q.hi[0] > c.lo[0]) & (q.lo[0] < c.hi[0]) & (q.hi[1] > c.lo[1]) & (q.lo[1] < c.hi[1]) &
           (q.hi[2] > c.lo[2]) & (q.lo[2] < c.hi[2]);
}

// One step of the DFS of collision.cl:174-226 at internal node `idx`: tests both children
// conservatively on the grid and returns the next node (TRV_DONE when the stack ran empty).
// Leaf children that pass are reported through on_leaf(leaf position), child a before child b:
//   EXACT  -- only after the exact test on the float boxes (a report is a collision row);
//   !EXACT -- as CANDIDATES, without touching the Bound array: the walk then has no
//             data-dependent side branch at all and the exact test runs later, batched.
template <bool EXACT, typename T, typename F>
__device__ __forceinline__ uint32_t visit(const unsigned char* __restrict__ wide, const T* __restrict__ bounds,
                                          const Query<T>& query, uint32_t q, uint32_t leaf_start, uint32_t idx,
                                          uint32_t* stack, int& sp, F&& on_leaf) {
    const QNode w = load_qnode(wide, idx);
    const uint32_t ca = w.w[0], cb = w.w[1];
    const bool la = (int32_t)ca < 0, lb = (int32_t)cb < 0;   // leaf: 0x80000000 | sorted position
    const uint32_t pa = ca & 0x7fffffffu, pb = cb & 0x7fffffffu;
    // child a is a left child: its right edge is its own index (leaf: its position);
    // child b's right edge is this node's, already known to be > q
    const bool oa = grid_overlap(query.grid, w.w[2], w.w[3], w.w[4]) && pa > q;
    const bool ob = grid_overlap(query.grid, w.w[5], w.w[6], w.w[7]);
    if (EXACT) {
        if (oa && la) {
            if (exact_overlap(query, load_leaf_box(bounds, leaf_start + pa))) on_leaf(pa, true);
        }
        if (ob && lb) {
            if (exact_overlap(query, load_leaf_box(bounds, leaf_start + pb))) on_leaf(pb, true);
        }
    } else {   // straight-line: the reports are predicated, not branched around
        on_leaf(pa, oa && la);
        on_leaf(pb, ob && lb);
    }
    const bool ta = oa && !la, tb = ob && !lb;
    if (!ta && !tb) return stack[--sp];
    if (ta && tb) {
#ifdef COL_DEBUG   // collision.cl:186 has the same unguarded 64-entry stack; a debug build traps instead of overrunning
        if (sp >= 64) __trap();
#endif
        stack[sp++] = cb;
    }
    return ta ? ca : cb;
}

// One whole walk of one query.  on_leaf(leaf position, k) for the k-th report, in
// reference order.  Returns the number of reports.
template <bool EXACT, typename T, typename F>
__device__ __forceinline__ uint32_t walk(const unsigned char* __restrict__ wide, const T* __restrict__ bounds,
                                         const Query<T>& query, uint32_t q, uint32_t leaf_start, F&& on_leaf) {
    if (q >= leaf_start) return 0;  // the last sorted leaf has nothing to its right (every right_edge <= q)
    uint32_t stack[64];
    int sp = 0;
    stack[sp++] = TRV_DONE;
    uint32_t idx = 0, hits = 0;
#ifdef COL_DEBUG
    uint32_t visits = 0;
#endif
    do {
#ifdef COL_DEBUG   // a walk over half-built records need not end: trap instead of hanging
        if (++visits > 2u * leaf_start + 2u || idx >= leaf_start) __trap();
#endif
        idx = visit<EXACT>(wide, bounds, query, q, leaf_start, idx, stack, sp, [&](uint32_t pos, bool valid) {
            on_leaf(pos, hits, valid);
            hits += valid ? 1u : 0u;
        });
    } while (idx != TRV_DONE);
    return hits;
}

__device__ __forceinline__ void block_add(unsigned long long* total, unsigned long long v) {
    __shared__ unsigned long long s_sum[TRV_THREADS / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane_id() == 0) s_sum[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0 && total) {
        unsigned long long t = 0;
        for (int w = 0; w < TRV_THREADS / 32; w++) t += s_sum[w];
        if (t) atomicAdd(total, t);
    }
}

// ---- count only ---------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(TRV_THREADS)
count_kernel(unsigned long long* __restrict__ total, uint32_t* __restrict__ counts,
             const unsigned char* __restrict__ wide, const T* __restrict__ bounds, const T* __restrict__ scene,
             uint32_t n, uint32_t q_begin, uint32_t q_end) {
    const uint32_t q = q_begin + blockIdx.x * TRV_THREADS + threadIdx.x;
    uint32_t hits = 0;
    if (q < q_end) {
        const uint32_t leaf_start = n - 1;
        const Query<T> query = load_query(bounds, leaf_start + q, make_quant(scene));
        hits = walk<true>(wide, bounds, query, q, leaf_start, [](uint32_t, uint32_t, bool) {});
        if (counts) counts[q - q_begin] = hits;
    }
    block_add(total, hits);
}

// ---- walk once: candidates ------------------------------------------------------
// Every query walks ONCE, on the traversal records only.  The leaves whose grid boxes
// overlap the query's are CANDIDATES: their number goes to counts[], the first `depth`
// of them (sorted leaf positions, in reference order) are parked in a k-major global
// array stash[k * nq + i]: lanes of a warp park their k-th candidate in one 128-byte row,
// and only rows that are actually used ever move.  No inter-block dependency, so dense
// and sparse regions of the scene do not stall each other.
template <typename T>
__global__ void __launch_bounds__(TRV_THREADS)
candidates_kernel(uint32_t* __restrict__ counts, uint32_t* __restrict__ stash, const unsigned char* __restrict__ wide,
                  const T* __restrict__ bounds, const T* __restrict__ scene, uint32_t n, uint32_t q_begin,
                  uint32_t q_end, uint32_t prefetch_lines, uint32_t depth) {
    const uint32_t nq = q_end - q_begin;
    const uint32_t i = blockIdx.x * TRV_THREADS + threadIdx.x;
    {   // the low levels of the walks of this block live in the records right of its first query
        const uint32_t first = q_begin + blockIdx.x * TRV_THREADS;
        const uint32_t rec = first + threadIdx.x * 4;  // one 128-byte line (4 records) per thread
        if (threadIdx.x < prefetch_lines && rec < n - 1)
            asm volatile("prefetch.global.L1 [%0];" ::"l"(wide + (size_t)rec * QNODE_BYTES));
    }
    if (i >= nq) return;
    const uint32_t q = q_begin + i, leaf_start = n - 1;
    const Query<T> query = load_query(bounds, leaf_start + q, make_quant(scene));
    uint32_t* park = stash + i;
    counts[i] = walk<false>(wide, bounds, query, q, leaf_start, [park, nq, depth](uint32_t pos, uint32_t k, bool valid) {
        if (valid && k < depth) park[(size_t)k * nq] = pos;
    });
}

// ---- confirm: candidates -> collision rows ----------------------------------------
// The exact test of collision.cl:164-166 on the float boxes, batched: thread i runs over
// its parked candidates (coalesced k-major rows), keeps the confirmed ones in place (same
// order) and replaces counts[i] by the number of rows of the query.  A query with more
// than `depth` candidates walks again, exactly, parking its first `depth` rows.
// What is kept is the partner's ID, not its position: ids[position] is fetched beside the leaf
// box (both depend on the position only, so the gather adds no link to the dependent chain here),
// and the emit kernel, whose chain was stash -> ids -> strip, copies what it finds.
template <typename T>
__global__ void __launch_bounds__(TRV_THREADS)
confirm_kernel(unsigned long long* __restrict__ total, uint32_t* __restrict__ counts, uint32_t* __restrict__ stash,
               const unsigned char* __restrict__ wide, const T* __restrict__ bounds, const T* __restrict__ scene,
               const uint32_t* __restrict__ ids, uint32_t n, uint32_t q_begin, uint32_t q_end, uint32_t depth) {
    const uint32_t nq = q_end - q_begin;
    const uint32_t i = blockIdx.x * TRV_THREADS + threadIdx.x;
    uint32_t hits = 0;
    if (i < nq) {
        const uint32_t cand = counts[i];
        if (cand) {
            const uint32_t q = q_begin + i, leaf_start = n - 1;
            uint32_t* park = stash + i;
            if (cand <= depth) {
                const LeafBox<T> b = load_leaf_box(bounds, leaf_start + q);
                Query<T> query;
#pragma unroll
                for (int k = 0; k < 3; k++) { query.lo[k] = b.lo[k]; query.hi[k] = b.hi[k]; }
                // software pipeline: position k+2 and box k+1 are in flight while box k is tested (the
                // compaction writes slot hits <= k, behind every slot still to be read)
                uint32_t p1 = cand > 1 ? park[nq] : 0u;
                uint32_t id0;
                LeafBox<T> b0;
                {
                    const uint32_t p0 = park[0];
                    b0 = load_leaf_box(bounds, leaf_start + p0);
                    id0 = __ldg(ids + p0);
                }
                for (uint32_t k = 0; k < cand; k++) {
                    LeafBox<T> b1 = b0;
                    uint32_t p2 = 0, id1 = 0;
                    if (k + 1 < cand) {
                        b1 = load_leaf_box(bounds, leaf_start + p1);
                        id1 = __ldg(ids + p1);
                    }
                    if (k + 2 < cand) p2 = park[(size_t)(k + 2) * nq];
                    if (exact_overlap(query, b0)) {
                        park[(size_t)hits * nq] = id0;
                        hits++;
                    }
                    id0 = id1;
                    p1 = p2;
                    b0 = b1;
                }
            } else {
                const Query<T> query = load_query(bounds, leaf_start + q, make_quant(scene));
                hits = walk<true>(wide, bounds, query, q, leaf_start, [park, nq, depth, ids](uint32_t pos, uint32_t k, bool) {
                    if (k < depth) park[(size_t)k * nq] = __ldg(ids + pos);
                });
            }
            counts[i] = hits;
        }
    }
    block_add(total, hits);
}

// ---- emit: rows to their final, deterministic positions -----------------------
// offsets = exclusive scan of counts.  Row k of query i lands at base + offsets[i] + k
// (query-major, DFS order: the order of the reference's work-items run sequentially).
// Parked hits are simply copied (ids gathered); a query with more than `depth` hits
// walks again for the remainder only.
#ifndef COL_EMIT_CHUNK
#define COL_EMIT_CHUNK 8
#endif
constexpr uint32_t EMIT_CHUNK = COL_EMIT_CHUNK;   // parked rows per query moved per round

template <typename T>
__global__ void __launch_bounds__(TRV_THREADS)
emit_kernel(uint32_t* __restrict__ pairs, uint64_t capacity, uint64_t base, const uint32_t* __restrict__ offsets,
            const unsigned long long* __restrict__ total, const uint32_t* __restrict__ stash,
            const unsigned char* __restrict__ wide, const T* __restrict__ bounds, const T* __restrict__ scene,
            const uint32_t* __restrict__ ids, uint32_t n, uint32_t q_begin, uint32_t q_end, uint32_t depth) {
    // A warp owns 32 consecutive queries, whose rows are (nearly) one contiguous piece of the
    // list.  Per round every lane gathers up to EMIT_CHUNK of its parked rows (coalesced
    // k-major reads) into the warp's shared-memory strip IN LIST ORDER; the strip then goes
    // out with full-width stores instead of one 8-byte store per lane into 32 different sectors.
    __shared__ uint32_t s_other[TRV_THREADS / 32][32 * EMIT_CHUNK];
    __shared__ unsigned char s_owner[TRV_THREADS / 32][32 * EMIT_CHUNK];
    const uint32_t i = blockIdx.x * TRV_THREADS + threadIdx.x;
    const uint32_t nq = q_end - q_begin;
    const unsigned warp = threadIdx.x >> 5, lane = lane_id();
    const bool live = i < nq;
    const uint32_t q = q_begin + i;
    uint32_t off = 0, hits = 0, qid = 0;
    if (live) {
        off = offsets[i];
        hits = ((i + 1 < nq) ? offsets[i + 1] : (uint32_t)*total) - off;
        if (base + off >= capacity) hits = 0;
        qid = __ldg(ids + q);
    }
    const uint32_t parked = min(hits, depth);
    const uint32_t most = __reduce_max_sync(0xffffffffu, parked);
    for (uint32_t k0 = 0; k0 < most; k0 += EMIT_CHUNK) {
        const uint32_t mine = parked > k0 ? min(parked - k0, EMIT_CHUNK) : 0u;   // my rows this round
        uint32_t start = mine;                                                     // -> exclusive prefix over the lanes
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, start, o);
            if (lane >= (unsigned)o) start += up;
        }
        const uint32_t rows = __shfl_sync(0xffffffffu, start, 31);
        start -= mine;
        {   // all loads of the round in flight together (the stash holds partner ids: see confirm_kernel)
            uint32_t other[EMIT_CHUNK];
#pragma unroll
            for (uint32_t k = 0; k < EMIT_CHUNK; k++)
                if (k < mine) other[k] = __ldg(stash + (size_t)(k0 + k) * nq + i);
#pragma unroll
            for (uint32_t k = 0; k < EMIT_CHUNK; k++)
                if (k < mine) {
                    s_other[warp][start + k] = other[k];
                    s_owner[warp][start + k] = (unsigned char)lane;
                }
        }
        __syncwarp();
        for (uint32_t f0 = 0; f0 < rows; f0 += 32) {
            const uint32_t f = f0 + lane;
            const bool on = f < rows;
            const unsigned owner = on ? s_owner[warp][f] : 0u;
            const uint32_t o_start = __shfl_sync(0xffffffffu, start, owner);
            const uint32_t o_off = __shfl_sync(0xffffffffu, off, owner);
            const uint32_t o_qid = __shfl_sync(0xffffffffu, qid, owner);
            const uint64_t at = base + o_off + k0 + (f - o_start);
            if (on && at < capacity) *reinterpret_cast<uint2*>(pairs + at * 2) = make_uint2(o_qid, s_other[warp][f]);
        }
        __syncwarp();
    }
    if (hits > depth) {   // the rows beyond the stash: walk again, exactly
        const uint32_t leaf_start = n - 1;
        const uint64_t out = base + off;
        const Query<T> query = load_query(bounds, leaf_start + q, make_quant(scene));
        walk<true>(wide, bounds, query, q, leaf_start, [=](uint32_t pos, uint32_t k, bool) {
            const uint64_t at = out + k;
            if (k >= depth && at < capacity)
                *reinterpret_cast<uint2*>(pairs + at * 2) = make_uint2(qid, __ldg(ids + pos));
        });
    }
}

// ---- write pass with given offsets (multi-GPU / external scan) --------------
template <typename T>
__global__ void __launch_bounds__(TRV_THREADS)
write_kernel(uint32_t* __restrict__ pairs, uint64_t capacity, const uint32_t* __restrict__ offsets, uint64_t base,
             const unsigned char* __restrict__ wide, const T* __restrict__ bounds, const T* __restrict__ scene,
             const uint32_t* __restrict__ ids, uint32_t n, uint32_t q_begin, uint32_t q_end) {
    const uint32_t q = q_begin + blockIdx.x * TRV_THREADS + threadIdx.x;
    if (q >= q_end) return;
    const uint32_t leaf_start = n - 1;
    const uint64_t out = base + offsets[q - q_begin];
    if (out >= capacity) return;
    const Query<T> query = load_query(bounds, leaf_start + q, make_quant(scene));
    const uint32_t qid = __ldg(ids + q);
    walk<true>(wide, bounds, query, q, leaf_start, [=](uint32_t pos, uint32_t k, bool) {
        const uint64_t at = out + k;
        if (at < capacity) *reinterpret_cast<uint2*>(pairs + at * 2) = make_uint2(qid, __ldg(ids + pos));
    });
}

inline bool bad_range(uint32_t n, uint32_t q_begin, uint32_t q_end) { return q_end > n || q_begin > q_end; }

}  // namespace

// Depth of the per-query stash: deep enough that few queries overflow it (an overflowing
// query walks two more times).  Chosen from the rows per query the caller made room for
// (normally the result of a count-only call); the same rule sizes the scratch.
static uint32_t stash_depth(uint32_t nq, uint64_t capacity) {
    const uint64_t per_query = capacity / (nq ? nq : 1);
    return per_query <= 6 ? TRV_STASH_MIN : per_query <= 24 ? 64u : TRV_STASH_MAX;
}

// scratch of col_traverse_wide_pairs: [counts/offsets: nq u32][scan scratch][stash: depth x nq u32]
struct PairsLayout { size_t counts, scan, stash, total; };
static PairsLayout pairs_layout(uint32_t nq, uint32_t depth) {
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    PairsLayout L;
    L.counts = 0;
    L.scan = up((size_t)nq * 4 + 4);
    L.stash = L.scan + up(col_scan_scratch_bytes(nq));
    L.total = L.stash + up((size_t)nq * 4 * depth);
    return L;
}

COL_API size_t col_traverse_scratch_bytes(uint32_t n_queries, uint64_t capacity) {
    const uint32_t nq = n_queries ? n_queries : 1;
    return pairs_layout(nq, stash_depth(nq, capacity)).total;
}

COL_API int col_traverse_wide_count(uint64_t* total, uint32_t* counts, const void* wide, const void* bounds,
                                    const void* scene2x4, int dtype, uint32_t n, uint32_t q_begin, uint32_t q_end,
                                    void* stream) {
    if (!wide || !bounds || !scene2x4 || bad_range(n, q_begin, q_end)) return COL_ERR_BAD_ARG;
    if (dtype != COL_F32 && dtype != COL_F64) return COL_ERR_UNSUPPORTED;
    cudaStream_t s = col_stream(stream);
    if (total) COL_CUDA(cudaMemsetAsync(total, 0, sizeof(uint64_t), s));
    if (q_begin == q_end) return 0;
    if (n < 2) return counts ? (int)cudaMemsetAsync(counts, 0, sizeof(uint32_t) * (q_end - q_begin), s) : 0;
    const unsigned blocks = col_blocks(q_end - q_begin, TRV_THREADS);
    const unsigned char* w = (const unsigned char*)wide;
    if (dtype == COL_F32)
        count_kernel<float><<<blocks, TRV_THREADS, 0, s>>>((unsigned long long*)total, counts, w, (const float*)bounds, (const float*)scene2x4, n, q_begin, q_end);
    else
        count_kernel<double><<<blocks, TRV_THREADS, 0, s>>>((unsigned long long*)total, counts, w, (const double*)bounds, (const double*)scene2x4, n, q_begin, q_end);
    COL_CHECK_LAUNCH();
    return 0;
}

template <typename T>
static int pairs_impl(uint32_t* pairs, uint64_t capacity, uint64_t base, unsigned long long* total,
                      const unsigned char* w, const T* bounds, const T* scene, const uint32_t* ids, uint32_t n,
                      uint32_t q_begin, uint32_t q_end, char* sc, void* stream, void* const* marks) {
    cudaStream_t s = col_stream(stream);
    auto mark = [&](int k) -> int {   // optional stage marks (cudaEvent_t handles) for the bench's compaction figure
        return (marks && marks[k]) ? (int)cudaEventRecord((cudaEvent_t)marks[k], s) : 0;
    };
    const uint32_t nq = q_end - q_begin;
    const uint32_t depth = stash_depth(nq, capacity);
    const PairsLayout L = pairs_layout(nq, depth);
    uint32_t* counts = reinterpret_cast<uint32_t*>(sc + L.counts);
    uint32_t* stash = reinterpret_cast<uint32_t*>(sc + L.stash);
    const unsigned blocks = col_blocks(nq, TRV_THREADS);
    const uint32_t pf = 64;   // 128-byte lines of the block's own record window prefetched to L1 (0..256: +4 % at 64)
    // (Round 2, measured and dropped -- profiles/r2e_traversal_variants.txt: walk + confirmation fused into
    //  one launch, 7.84 vs 7.48 ms at configs[3]; four queries per lane in the walk, 12.4 ms; four leaf boxes
    //  in flight per lane in the confirmation, 2.25 vs 1.99 ms; an emit that builds whole rows per window of
    //  the warp's strip, 2.13 vs 1.70 ms.)
    candidates_kernel<T><<<blocks, TRV_THREADS, 0, s>>>(counts, stash, w, bounds, scene, n, q_begin, q_end, pf, depth);
    COL_CHECK_LAUNCH();
    COL_CUDA((cudaError_t)mark(0));
    confirm_kernel<T><<<blocks, TRV_THREADS, 0, s>>>(total, counts, stash, w, bounds, scene, ids, n, q_begin, q_end, depth);
    COL_CHECK_LAUNCH();
    COL_CUDA((cudaError_t)mark(1));
    const int rc = col_exclusive_scan_u32(counts, nq, nullptr, sc + L.scan, stream);
    if (rc) return rc;
    COL_CUDA((cudaError_t)mark(2));
    emit_kernel<T><<<blocks, TRV_THREADS, 0, s>>>(pairs, capacity, base, counts, total, stash, w, bounds, scene, ids, n, q_begin, q_end, depth);
    COL_CHECK_LAUNCH();
    return 0;
}

static int wide_pairs(uint32_t* pairs, uint64_t capacity, uint64_t base, uint64_t* total_out,
                      const void* wide, const void* bounds, const void* scene2x4,
                      const uint32_t* ids_sorted, int dtype, uint32_t n, uint32_t q_begin,
                      uint32_t q_end, void* scratch, void* stream, void* const* marks) {
    if (!wide || !bounds || !scene2x4 || !ids_sorted || !scratch || !total_out || bad_range(n, q_begin, q_end) ||
        (!pairs && capacity))
        return COL_ERR_BAD_ARG;
    if (dtype != COL_F32 && dtype != COL_F64) return COL_ERR_UNSUPPORTED;
    if (capacity == 0)
        return col_traverse_wide_count(total_out, nullptr, wide, bounds, scene2x4, dtype, n, q_begin, q_end, stream);
    COL_CUDA(cudaMemsetAsync(total_out, 0, sizeof(uint64_t), col_stream(stream)));
    if (q_begin == q_end || n < 2) return 0;
    unsigned long long* total = reinterpret_cast<unsigned long long*>(total_out);
    const unsigned char* w = (const unsigned char*)wide;
    char* sc = reinterpret_cast<char*>(scratch);
    if (dtype == COL_F32)
        return pairs_impl<float>(pairs, capacity, base, total, w, (const float*)bounds, (const float*)scene2x4, ids_sorted, n, q_begin, q_end, sc, stream, marks);
    return pairs_impl<double>(pairs, capacity, base, total, w, (const double*)bounds, (const double*)scene2x4, ids_sorted, n, q_begin, q_end, sc, stream, marks);
}

COL_API int col_traverse_wide_pairs(uint32_t* pairs, uint64_t capacity, uint64_t base, uint64_t* total_out,
                                    const void* wide, const void* bounds, const void* scene2x4,
                                    const uint32_t* ids_sorted, int dtype, uint32_t n, uint32_t q_begin,
                                    uint32_t q_end, void* scratch, void* stream) {
    return wide_pairs(pairs, capacity, base, total_out, wide, bounds, scene2x4, ids_sorted, dtype, n, q_begin, q_end,
                      scratch, stream, nullptr);
}

COL_API int col_traverse_wide_pairs_marked(uint32_t* pairs, uint64_t capacity, uint64_t base, uint64_t* total_out,
                                           const void* wide, const void* bounds, const void* scene2x4,
                                           const uint32_t* ids_sorted, int dtype, uint32_t n, uint32_t q_begin,
                                           uint32_t q_end, void* scratch, void* stream, void* const* h_marks3) {
    return wide_pairs(pairs, capacity, base, total_out, wide, bounds, scene2x4, ids_sorted, dtype, n, q_begin, q_end,
                      scratch, stream, h_marks3);
}

COL_API int col_traverse_wide_write(uint32_t* pairs, uint64_t capacity, const uint32_t* offsets, uint64_t base,
                                    const void* wide, const void* bounds, const void* scene2x4,
                                    const uint32_t* ids_sorted, int dtype, uint32_t n, uint32_t q_begin,
                                    uint32_t q_end, void* stream) {
    if (!wide || !bounds || !scene2x4 || !ids_sorted || !offsets || bad_range(n, q_begin, q_end) || (!pairs && capacity))
        return COL_ERR_BAD_ARG;
    if (dtype != COL_F32 && dtype != COL_F64) return COL_ERR_UNSUPPORTED;
    if (capacity == 0 || q_begin == q_end || n < 2) return 0;
    const unsigned blocks = col_blocks(q_end - q_begin, TRV_THREADS);
    cudaStream_t s = col_stream(stream);
    const unsigned char* w = (const unsigned char*)wide;
    if (dtype == COL_F32)
        write_kernel<float><<<blocks, TRV_THREADS, 0, s>>>(pairs, capacity, offsets, base, w, (const float*)bounds, (const float*)scene2x4, ids_sorted, n, q_begin, q_end);
    else
        write_kernel<double><<<blocks, TRV_THREADS, 0, s>>>(pairs, capacity, offsets, base, w, (const double*)bounds, (const double*)scene2x4, ids_sorted, n, q_begin, q_end);
    COL_CHECK_LAUNCH();
    return 0;
}
