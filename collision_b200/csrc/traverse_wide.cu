// traverse_wide.cu -- the Collider's traversal: wide nodes + single-pass
// deterministic pair compaction.
//
// Same walk as collision/collision.cl:174-226 (see traverse.cu, which keeps the
// reference's Node/Bound argument contract), restructured around HBM/L1 traffic:
//
//  * one 64-byte (f32) / 128-byte (f64) record per internal node holds both
//    children's boxes, indices and right edges (written by tree.cu + refit.cu):
//    a visit is two 256-bit loads from one line instead of seven 128-bit loads
//    from five lines (the profile of the Node/Bound walk is L1-sector bound).
//  * count -> scan -> write without a second walk: every query walks once and
//    parks its first TRV_STASH hits in a k-major global array (coalesced rows, only
//    the used ones move); the counts are scanned by the single-pass scan kernel;
//    an emit kernel copies the parked hits to their final rows.  The row order is
//    exactly that of the reference's work-items run sequentially (query-major,
//    DFS order, child a before b).  Only queries with more than TRV_STASH hits walk
//    a second time, for the remainder.  (A single fused kernel chaining blocks by
//    decoupled look-back was measured slower: blocks of dense regions stall every
//    later block, and a shared-memory stash shrinks the L1 the walk depends on.)
#include "common.cuh"

namespace {

constexpr int TRV_THREADS = 128;
constexpr int TRV_STASH = 32;

template <typename T> struct WNode {
    T a[6], b[6];
    uint32_t ca, cb, ra, rb;
};

__device__ __forceinline__ void load_wide(const float* __restrict__ wide, uint32_t idx, WNode<float>& w) {
    const float* p = wide + (size_t)idx * 16;
    float v[8];
    uint32_t u[4];
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(w.a[0]), "=f"(w.a[1]), "=f"(w.a[2]), "=f"(w.a[3]), "=f"(w.a[4]), "=f"(w.a[5]), "=f"(w.b[0]),
                   "=f"(w.b[1])
                 : "l"(p));
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3])
                 : "l"(p + 8));
    w.b[2] = v[0]; w.b[3] = v[1]; w.b[4] = v[2]; w.b[5] = v[3];
    w.ca = u[0]; w.cb = u[1]; w.ra = u[2]; w.rb = u[3];
}
__device__ __forceinline__ void load_wide(const double* __restrict__ wide, uint32_t idx, WNode<double>& w) {
    const double* p = wide + (size_t)idx * 16;
    double v[12];
#pragma unroll
    for (int k = 0; k < 3; k++)
        asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];"
                     : "=d"(v[4 * k]), "=d"(v[4 * k + 1]), "=d"(v[4 * k + 2]), "=d"(v[4 * k + 3])
                     : "l"(p + 4 * k));
    const uint4 m = __ldg(reinterpret_cast<const uint4*>(p + 12));
#pragma unroll
    for (int k = 0; k < 6; k++) { w.a[k] = v[k]; w.b[k] = v[6 + k]; }
    w.ca = m.x; w.cb = m.y; w.ra = m.z; w.rb = m.w;
}

template <typename T> struct Query { T lo[3], hi[3]; };

__device__ __forceinline__ Query<float> load_query(const float* __restrict__ bounds, uint32_t node) {
    const float4* p = reinterpret_cast<const float4*>(bounds) + (size_t)node * 2;
    const float4 a = __ldg(p), c = __ldg(p + 1);
    return {{a.x, a.y, a.z}, {c.x, c.y, c.z}};
}
__device__ __forceinline__ Query<double> load_query(const double* __restrict__ bounds, uint32_t node) {
    const double2* p = reinterpret_cast<const double2*>(bounds) + (size_t)node * 4;
    const double2 a = __ldg(p), c = __ldg(p + 1), d = __ldg(p + 2), e = __ldg(p + 3);
    return {{a.x, a.y, c.x}, {d.x, d.y, e.x}};
}

// collision.cl:164-166: all(q.max > c.min & q.min < c.max), strict, x y z
template <typename T>
__device__ __forceinline__ bool overlap(const Query<T>& q, const T (&c)[6]) {
    return (q.hi[0] > c[0]) & (q.lo[0] < c[3]) & (q.hi[1] > c[1]) & (q.lo[1] < c[4]) & (q.hi[2] > c[2]) &
           (q.lo[2] < c[5]);
}

// One DFS walk.  on_hit(leaf_position) is called for every reported leaf in
// reference order.  Returns the number of hits.
template <typename T, typename F>
__device__ __forceinline__ uint32_t walk(const T* __restrict__ wide, const Query<T>& query, uint32_t q,
                                         uint32_t leaf_start, F&& on_hit) {
    uint32_t stack[64];
    int sp = 0;
    stack[sp++] = 0xFFFFFFFFu;
    uint32_t idx = 0, hits = 0;
    do {
        WNode<T> w;
        load_wide(wide, idx, w);
        const bool oa = overlap(query, w.a) && !(w.ra <= q);
        const bool ob = overlap(query, w.b) && !(w.rb <= q);
        const bool la = w.ca >= leaf_start, lb = w.cb >= leaf_start;
        if (oa && la) { on_hit(w.ca - leaf_start, hits); hits++; }
        if (ob && lb) { on_hit(w.cb - leaf_start, hits); hits++; }
        const bool ta = oa && !la, tb = ob && !lb;
        if (!ta && !tb) {
            idx = stack[--sp];
        } else {
            idx = ta ? w.ca : w.cb;
            if (ta && tb) stack[sp++] = w.cb;
        }
    } while (idx != 0xFFFFFFFFu);
    return hits;
}

// ---- count only: per-block sum -> one atomic per block ----------------------
template <typename T>
__global__ void __launch_bounds__(TRV_THREADS)
count_kernel(unsigned long long* __restrict__ total, uint32_t* __restrict__ counts, const T* __restrict__ wide,
             const T* __restrict__ bounds, uint32_t n, uint32_t q_begin, uint32_t q_end) {
    const uint32_t q = q_begin + blockIdx.x * TRV_THREADS + threadIdx.x;
    const uint32_t leaf_start = n - 1;
    uint32_t hits = 0;
    if (q < q_end) {
        const Query<T> query = load_query(bounds, leaf_start + q);
        hits = walk(wide, query, q, leaf_start, [](uint32_t, uint32_t) {});
        if (counts) counts[q - q_begin] = hits;
    }
    __shared__ uint32_t s_sum[TRV_THREADS / 32];
    uint32_t v = hits;
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

// ---- walk once: per-query counts + parked hits ---------------------------------
// Every query walks ONCE.  Its hit count goes to counts[], its first TRV_STASH hits
// (sorted leaf positions) are parked in a k-major global array stash[k * nq + i]:
// lanes of a warp park their k-th hit in one 128-byte row, and only rows that
// are actually used ever move.  No inter-block dependency, so dense and sparse
// regions of the scene do not stall each other.
template <typename T>
__global__ void __launch_bounds__(TRV_THREADS)
walk_park_kernel(unsigned long long* __restrict__ total, uint32_t* __restrict__ counts, uint32_t* __restrict__ stash,
                 const T* __restrict__ wide, const T* __restrict__ bounds, uint32_t n, uint32_t q_begin,
                 uint32_t q_end) {
    const uint32_t i = blockIdx.x * TRV_THREADS + threadIdx.x;
    const uint32_t q = q_begin + i;
    const uint32_t nq = q_end - q_begin;
    const uint32_t leaf_start = n - 1;
    uint32_t hits = 0;
    if (q < q_end) {
        const Query<T> query = load_query(bounds, leaf_start + q);
        uint32_t* park = stash + i;
        hits = walk(wide, query, q, leaf_start, [park, nq](uint32_t pos, uint32_t k) {
            if (k < TRV_STASH) park[(size_t)k * nq] = pos;
        });
        counts[i] = hits;
    }
    __shared__ uint32_t s_sum[TRV_THREADS / 32];
    uint32_t v = hits;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane_id() == 0) s_sum[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < TRV_THREADS / 32; w++) t += s_sum[w];
        if (t) atomicAdd(total, t);
    }
}

// ---- emit: rows to their final, deterministic positions -----------------------
// offsets = exclusive scan of counts.  Row k of query i lands at base + offsets[i] + k
// (query-major, DFS order: the order of the reference's work-items run sequentially).
// Parked hits are simply copied (ids gathered); a query with more than TRV_STASH hits
// walks again for the remainder only.
template <typename T>
__global__ void __launch_bounds__(TRV_THREADS)
emit_kernel(uint32_t* __restrict__ pairs, uint64_t capacity, uint64_t base, const uint32_t* __restrict__ offsets,
            const unsigned long long* __restrict__ total, const uint32_t* __restrict__ stash,
            const T* __restrict__ wide, const T* __restrict__ bounds, const uint32_t* __restrict__ ids, uint32_t n,
            uint32_t q_begin, uint32_t q_end) {
    const uint32_t i = blockIdx.x * TRV_THREADS + threadIdx.x;
    const uint32_t q = q_begin + i;
    if (q >= q_end) return;
    const uint32_t nq = q_end - q_begin;
    const uint32_t off = offsets[i];
    const uint32_t next = (i + 1 < nq) ? offsets[i + 1] : (uint32_t)*total;
    const uint32_t hits = next - off;
    const uint64_t out = base + off;
    if (hits == 0 || out >= capacity) return;
    const uint32_t qid = __ldg(ids + q);
    const uint32_t parked = min(hits, (uint32_t)TRV_STASH);
    for (uint32_t k = 0; k < parked; k++) {
        const uint64_t at = out + k;
        if (at < capacity)
            *reinterpret_cast<uint2*>(pairs + at * 2) = make_uint2(qid, __ldg(ids + __ldg(stash + (size_t)k * nq + i)));
    }
    if (hits > TRV_STASH) {
        const uint32_t leaf_start = n - 1;
        const Query<T> query = load_query(bounds, leaf_start + q);
        walk(wide, query, q, leaf_start, [=](uint32_t pos, uint32_t k) {
            const uint64_t at = out + k;
            if (k >= TRV_STASH && at < capacity)
                *reinterpret_cast<uint2*>(pairs + at * 2) = make_uint2(qid, __ldg(ids + pos));
        });
    }
}

// ---- write pass with given offsets (multi-GPU / external scan) --------------
template <typename T>
__global__ void __launch_bounds__(TRV_THREADS)
write_kernel(uint32_t* __restrict__ pairs, uint64_t capacity, const uint32_t* __restrict__ offsets, uint64_t base,
             const T* __restrict__ wide, const T* __restrict__ bounds, const uint32_t* __restrict__ ids, uint32_t n,
             uint32_t q_begin, uint32_t q_end) {
    const uint32_t q = q_begin + blockIdx.x * TRV_THREADS + threadIdx.x;
    if (q >= q_end) return;
    const uint32_t leaf_start = n - 1;
    const uint64_t out = base + offsets[q - q_begin];
    if (out >= capacity) return;
    const Query<T> query = load_query(bounds, leaf_start + q);
    const uint32_t qid = __ldg(ids + q);
    walk(wide, query, q, leaf_start, [=](uint32_t pos, uint32_t k) {
        const uint64_t at = out + k;
        if (at < capacity) *reinterpret_cast<uint2*>(pairs + at * 2) = make_uint2(qid, __ldg(ids + pos));
    });
}

inline bool bad_range(uint32_t n, uint32_t q_begin, uint32_t q_end) { return q_end > n || q_begin > q_end; }

}  // namespace

// scratch of col_traverse_wide_pairs: [counts/offsets: nq u32][scan scratch][stash: TRV_STASH x nq u32]
struct PairsLayout { size_t counts, scan, stash, total; };
static PairsLayout pairs_layout(uint32_t nq) {
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    PairsLayout L;
    L.counts = 0;
    L.scan = up((size_t)nq * 4 + 4);
    L.stash = L.scan + up(col_scan_scratch_bytes(nq));
    L.total = L.stash + up((size_t)nq * 4 * TRV_STASH);
    return L;
}

COL_API size_t col_traverse_scratch_bytes(uint32_t n_queries) { return pairs_layout(n_queries ? n_queries : 1).total; }

COL_API int col_traverse_wide_count(uint64_t* total, uint32_t* counts, const void* wide, const void* bounds, int dtype,
                                    uint32_t n, uint32_t q_begin, uint32_t q_end, void* stream) {
    if (!wide || !bounds || bad_range(n, q_begin, q_end)) return COL_ERR_BAD_ARG;
    cudaStream_t s = col_stream(stream);
    if (total) COL_CUDA(cudaMemsetAsync(total, 0, sizeof(uint64_t), s));
    if (q_begin == q_end) return 0;
    if (n < 2) return counts ? (int)cudaMemsetAsync(counts, 0, sizeof(uint32_t) * (q_end - q_begin), s) : 0;
    const unsigned blocks = col_blocks(q_end - q_begin, TRV_THREADS);
    if (dtype == COL_F32)
        count_kernel<float><<<blocks, TRV_THREADS, 0, s>>>((unsigned long long*)total, counts, (const float*)wide, (const float*)bounds, n, q_begin, q_end);
    else if (dtype == COL_F64)
        count_kernel<double><<<blocks, TRV_THREADS, 0, s>>>((unsigned long long*)total, counts, (const double*)wide, (const double*)bounds, n, q_begin, q_end);
    else
        return COL_ERR_UNSUPPORTED;
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_traverse_wide_pairs(uint32_t* pairs, uint64_t capacity, uint64_t base, uint64_t* total_out,
                                    const void* wide, const void* bounds, const uint32_t* ids_sorted, int dtype,
                                    uint32_t n, uint32_t q_begin, uint32_t q_end, void* scratch, void* stream) {
    if (!wide || !bounds || !ids_sorted || !scratch || !total_out || bad_range(n, q_begin, q_end) || (!pairs && capacity))
        return COL_ERR_BAD_ARG;
    if (dtype != COL_F32 && dtype != COL_F64) return COL_ERR_UNSUPPORTED;
    if (capacity == 0) return col_traverse_wide_count(total_out, nullptr, wide, bounds, dtype, n, q_begin, q_end, stream);
    cudaStream_t s = col_stream(stream);
    COL_CUDA(cudaMemsetAsync(total_out, 0, sizeof(uint64_t), s));
    if (q_begin == q_end || n < 2) return 0;
    const uint32_t nq = q_end - q_begin;
    const PairsLayout L = pairs_layout(nq);
    char* sc = reinterpret_cast<char*>(scratch);
    uint32_t* counts = reinterpret_cast<uint32_t*>(sc + L.counts);
    uint32_t* stash = reinterpret_cast<uint32_t*>(sc + L.stash);
    unsigned long long* total = reinterpret_cast<unsigned long long*>(total_out);
    const unsigned blocks = col_blocks(nq, TRV_THREADS);
    if (dtype == COL_F32)
        walk_park_kernel<float><<<blocks, TRV_THREADS, 0, s>>>(total, counts, stash, (const float*)wide, (const float*)bounds, n, q_begin, q_end);
    else
        walk_park_kernel<double><<<blocks, TRV_THREADS, 0, s>>>(total, counts, stash, (const double*)wide, (const double*)bounds, n, q_begin, q_end);
    COL_CHECK_LAUNCH();
    const int rc = col_exclusive_scan_u32(counts, nq, nullptr, sc + L.scan, stream);
    if (rc) return rc;
    if (dtype == COL_F32)
        emit_kernel<float><<<blocks, TRV_THREADS, 0, s>>>(pairs, capacity, base, counts, total, stash, (const float*)wide, (const float*)bounds, ids_sorted, n, q_begin, q_end);
    else
        emit_kernel<double><<<blocks, TRV_THREADS, 0, s>>>(pairs, capacity, base, counts, total, stash, (const double*)wide, (const double*)bounds, ids_sorted, n, q_begin, q_end);
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_traverse_wide_write(uint32_t* pairs, uint64_t capacity, const uint32_t* offsets, uint64_t base,
                                    const void* wide, const void* bounds, const uint32_t* ids_sorted, int dtype,
                                    uint32_t n, uint32_t q_begin, uint32_t q_end, void* stream) {
    if (!wide || !bounds || !ids_sorted || !offsets || bad_range(n, q_begin, q_end) || (!pairs && capacity))
        return COL_ERR_BAD_ARG;
    if (capacity == 0 || q_begin == q_end || n < 2) return 0;
    const unsigned blocks = col_blocks(q_end - q_begin, TRV_THREADS);
    cudaStream_t s = col_stream(stream);
    if (dtype == COL_F32)
        write_kernel<float><<<blocks, TRV_THREADS, 0, s>>>(pairs, capacity, offsets, base, (const float*)wide, (const float*)bounds, ids_sorted, n, q_begin, q_end);
    else if (dtype == COL_F64)
        write_kernel<double><<<blocks, TRV_THREADS, 0, s>>>(pairs, capacity, offsets, base, (const double*)wide, (const double*)bounds, ids_sorted, n, q_begin, q_end);
    else
        return COL_ERR_UNSUPPORTED;
    COL_CHECK_LAUNCH();
    return 0;
}
