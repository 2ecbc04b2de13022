// traverse.cu -- stack-based BVH traversal, one thread per query leaf.
//
// Replaces traverse / checkOverlap / isLeaf (collision/collision.cl:164-226).
// The walk is the reference's, step for step: DFS from the root with a 64-entry
// private stack seeded with a sentinel, both children tested per visit with the
// strict per-axis box test (q.max > c.min && q.min < c.max on x, y, z; plain
// compares, nothing to contract), children whose right_edge <= query pruned
// (each unordered pair reported once, never a self pair), leaf hits emitted
// child a before child b.  What changes is how rows are committed:
//
//   count  : only the number of hits per query is stored
//   write  : rows go to offsets[q] + k  -> the exact order of the reference's
//            work-items run sequentially (query-major, DFS order)
//   atomic : the reference's own contract, one global counter, arbitrary
//            order (warp-aggregated), counter = true total beyond capacity
//
// count -> exclusive scan (scan.cu) -> write is the count/scan/offset
// compaction that makes the pair list deterministic.
#include "common.cuh"

namespace {

enum Mode { MODE_COUNT = 0, MODE_WRITE = 1, MODE_ATOMIC = 2 };

template <typename T> struct QBox { T lo[3], hi[3]; };

__device__ __forceinline__ QBox<float> load_box(const float* __restrict__ bounds, uint32_t node) {
    const float4* p = reinterpret_cast<const float4*>(bounds) + (size_t)node * 2;
    const float4 a = __ldg(p), c = __ldg(p + 1);
    QBox<float> b;
    b.lo[0] = a.x; b.lo[1] = a.y; b.lo[2] = a.z;
    b.hi[0] = c.x; b.hi[1] = c.y; b.hi[2] = c.z;
    return b;
}
__device__ __forceinline__ QBox<double> load_box(const double* __restrict__ bounds, uint32_t node) {
    const double2* p = reinterpret_cast<const double2*>(bounds) + (size_t)node * 4;
    const double2 a = __ldg(p), c = __ldg(p + 1), d = __ldg(p + 2), e = __ldg(p + 3);
    QBox<double> b;
    b.lo[0] = a.x; b.lo[1] = a.y; b.lo[2] = c.x;
    b.hi[0] = d.x; b.hi[1] = d.y; b.hi[2] = e.x;
    return b;
}

template <typename T>
__device__ __forceinline__ bool overlap(const QBox<T>& q, const QBox<T>& c) {
    return (q.hi[0] > c.lo[0]) & (q.lo[0] < c.hi[0]) & (q.hi[1] > c.lo[1]) & (q.lo[1] < c.hi[1]) &
           (q.hi[2] > c.lo[2]) & (q.lo[2] < c.hi[2]);
}

template <typename T, int MODE>
__global__ void __launch_bounds__(128)
traverse_kernel(uint32_t* __restrict__ pairs, uint32_t* __restrict__ counts_or_counter,
                const uint32_t* __restrict__ offsets, uint64_t base, uint64_t capacity,
                const col_node* __restrict__ nodes, const T* __restrict__ bounds, uint32_t n, uint32_t q_begin,
                uint32_t q_end) {
    const uint32_t q = q_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= q_end) return;
    const uint32_t leaf_start = n - 1;
    const QBox<T> query = load_box(bounds, leaf_start + q);
    const uint32_t qid = (MODE == MODE_COUNT) ? 0u : __ldg(&nodes[leaf_start + q].data[0]);
    uint64_t out = 0;
    if (MODE == MODE_WRITE) out = base + offsets[q - q_begin];
    uint32_t hits = 0;

    uint32_t stack[64];
    int sp = 0;
    stack[sp++] = 0xFFFFFFFFu;
    uint32_t idx = 0;
    do {
        const uint4 rec = __ldg(reinterpret_cast<const uint4*>(nodes + idx));
        const uint32_t ca = rec.z, cb = rec.w;
        const uint4 ra = __ldg(reinterpret_cast<const uint4*>(nodes + ca));
        const uint4 rb = __ldg(reinterpret_cast<const uint4*>(nodes + cb));
        const bool oa = overlap(query, load_box(bounds, ca)) && !(ra.y <= q);
        const bool ob = overlap(query, load_box(bounds, cb)) && !(rb.y <= q);
        const bool la = ca >= leaf_start, lb = cb >= leaf_start;

#pragma unroll
        for (int side = 0; side < 2; side++) {
            const bool hit = side == 0 ? (oa && la) : (ob && lb);
            const uint32_t other = side == 0 ? ra.z : rb.z;
            if (hit) {
                if (MODE == MODE_WRITE) {
                    const uint64_t at = out + hits;
                    if (at < capacity) *reinterpret_cast<uint2*>(pairs + at * 2) = make_uint2(qid, other);
                } else if (MODE == MODE_ATOMIC) {
                    const unsigned active = __activemask();
                    const int leader = __ffs(active) - 1;
                    uint32_t at = 0;
                    if ((int)lane_id() == leader) at = atomicAdd(counts_or_counter, (uint32_t)__popc(active));
                    at = __shfl_sync(active, at, leader) + __popc(active & lanemask_lt());
                    if (at < capacity) *reinterpret_cast<uint2*>(pairs + (size_t)at * 2) = make_uint2(qid, other);
                }
                hits++;
            }
        }
        const bool ta = oa && !la, tb = ob && !lb;
        if (!ta && !tb) {
            idx = stack[--sp];
        } else {
            idx = ta ? ca : cb;
            if (ta && tb) {
#ifdef COL_DEBUG   // (collision.cl:186: the reference's stack is unguarded too)
                if (sp >= 64) __trap();
#endif
                stack[sp++] = cb;
            }
        }
    } while (idx != 0xFFFFFFFFu);

    if (MODE == MODE_COUNT) counts_or_counter[q - q_begin] = hits;
}

template <int MODE>
int launch(uint32_t* pairs, uint32_t* cc, const uint32_t* offsets, uint64_t base, uint64_t capacity, const void* nodes,
           const void* bounds, int dtype, uint32_t n, uint32_t q_begin, uint32_t q_end, cudaStream_t s) {
    if (!nodes || !bounds || q_end > n || q_begin > q_end) return COL_ERR_BAD_ARG;
    if (q_begin == q_end) return 0;
    const unsigned blocks = col_blocks(q_end - q_begin, 128);
    if (n < 2) {  // a single sphere has no internal node: zero pairs (undefined upstream)
        if (MODE == MODE_COUNT) return (int)cudaMemsetAsync(cc, 0, sizeof(uint32_t) * (q_end - q_begin), s);
        return 0;
    }
    if (dtype == COL_F32)
        traverse_kernel<float, MODE><<<blocks, 128, 0, s>>>(pairs, cc, offsets, base, capacity, (const col_node*)nodes,
                                                            (const float*)bounds, n, q_begin, q_end);
    else if (dtype == COL_F64)
        traverse_kernel<double, MODE><<<blocks, 128, 0, s>>>(pairs, cc, offsets, base, capacity, (const col_node*)nodes,
                                                             (const double*)bounds, n, q_begin, q_end);
    else
        return COL_ERR_UNSUPPORTED;
    COL_CHECK_LAUNCH();
    return 0;
}

}  // namespace

COL_API int col_traverse_count(uint32_t* counts, const void* nodes, const void* bounds, int dtype, uint32_t n,
                               uint32_t q_begin, uint32_t q_end, void* stream) {
    if (!counts) return COL_ERR_BAD_ARG;
    return launch<MODE_COUNT>(nullptr, counts, nullptr, 0, 0, nodes, bounds, dtype, n, q_begin, q_end, col_stream(stream));
}

COL_API int col_traverse_write(uint32_t* pairs, uint64_t capacity, const uint32_t* offsets, uint64_t base,
                               const void* nodes, const void* bounds, int dtype, uint32_t n, uint32_t q_begin,
                               uint32_t q_end, void* stream) {
    if (!offsets || (!pairs && capacity)) return COL_ERR_BAD_ARG;
    if (capacity == 0) return 0;
    return launch<MODE_WRITE>(pairs, nullptr, offsets, base, capacity, nodes, bounds, dtype, n, q_begin, q_end,
                              col_stream(stream));
}

COL_API int col_traverse_atomic(uint32_t* pairs, uint32_t* counter, uint32_t capacity, const void* nodes,
                                const void* bounds, int dtype, uint32_t n, uint32_t q_begin, uint32_t q_end,
                                void* stream) {
    if (!counter || (!pairs && capacity)) return COL_ERR_BAD_ARG;
    return launch<MODE_ATOMIC>(pairs, counter, nullptr, 0, capacity, nodes, bounds, dtype, n, q_begin, q_end,
                               col_stream(stream));
}
