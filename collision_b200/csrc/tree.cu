// tree.cu -- Karras (2012) binary radix tree over the sorted Morton codes.
//
// Replaces fillInternal + generateBVH / delta (collision/collision.cl:55-121),
// which the reference launches as two kernels, with one launch: thread i < n
// writes leaf i (id, right_edge), thread i < n-1 also builds internal node i.
// The arithmetic follows the reference exactly -- same delta (common prefix
// of the codes, ties broken on the sorted position with 32 + clz(i ^ j), -1
// out of range), same doubling / binary searches, same child numbering -- so
// the Node array is bit-identical.  Every 16-byte record is written by two
// parties: {right_edge, data[2]} by its own thread and .parent by the parent's
// thread, exactly as upstream; the two fields upstream never writes
// (nodes[0].parent, leaf data[1]) are set to NO_NODE.
//
// Not bandwidth-bound: ~2*log2(range) dependent, cache-resident code loads per
// node; algorithmic traffic 8 B read + 32 B written per sphere.
#include "common.cuh"

namespace {

__device__ __forceinline__ int delta(const uint32_t* __restrict__ codes, uint32_t n, uint32_t i, uint32_t ci,
                                     bool forward, uint32_t offset) {
    if (forward && (i + offset) >= n) return -1;
    if (!forward && offset > i) return -1;
    const uint32_t j = forward ? i + offset : i - offset;
    const uint32_t cj = __ldg(codes + j);
    return ci != cj ? __clz(ci ^ cj) : 32 + __clz(i ^ j);
}

// WIDE: additionally write {child a, child b} into the 32-byte traversal record of i
// (qnode.cuh) and initialise the refit's arrival word.
template <bool LEAVES, bool WIDE>
__global__ void __launch_bounds__(256)
build_tree_kernel(const uint32_t* __restrict__ codes, const uint32_t* __restrict__ ids, col_node* __restrict__ nodes,
                  uint32_t n, unsigned char* __restrict__ wide, uint32_t* __restrict__ flags) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t leaf_start = n - 1;
    if (LEAVES) {
        // {right_edge, id, NO_NODE}: 12 bytes at byte offset 4 of the record
        uint32_t* rec = reinterpret_cast<uint32_t*>(nodes + leaf_start + i);
        rec[1] = i;
        rec[2] = __ldg(ids + i);
        rec[3] = 0xFFFFFFFFu;
        if (i == 0) nodes[0].parent = 0xFFFFFFFFu;
    }
    if (i >= leaf_start) return;

    const uint32_t ci = __ldg(codes + i);
    const bool forward = delta(codes, n, i, ci, true, 1) > delta(codes, n, i, ci, false, 1);
    const int delta_min = delta(codes, n, i, ci, !forward, 1);

    uint32_t len_max = 2;
    while (delta(codes, n, i, ci, forward, len_max) > delta_min) len_max *= 2;

    uint32_t len = 0;
    for (uint32_t t = len_max / 2; t > 0; t /= 2)
        if (delta(codes, n, i, ci, forward, len + t) > delta_min) len += t;

    const uint32_t j = forward ? i + len : i - len;
    const int delta_node = delta(codes, n, i, ci, forward, len) & 0xFF;  // unsigned char upstream
    uint32_t s = 0;
    {
        uint32_t t = len;
        do {
            t = (t + 1) / 2;
            if (delta(codes, n, i, ci, forward, s + t) > delta_node) s += t;
        } while (t > 1);
    }
    const uint32_t gamma = forward ? i + s : i - s - 1;
    const uint32_t lo = min(i, j), hi = max(i, j);
    const uint32_t child_a = (lo == gamma) ? leaf_start + gamma : gamma;
    const uint32_t child_b = (hi == gamma + 1) ? leaf_start + gamma + 1 : gamma + 1;

    uint32_t* rec = reinterpret_cast<uint32_t*>(nodes + i);
    rec[1] = hi;
    rec[2] = child_a;
    rec[3] = child_b;
    nodes[child_a].parent = i;
    nodes[child_b].parent = i;
    if (WIDE) {
        // children in the record: a leaf as 0x80000000 | its sorted position, an internal node as its index
        *reinterpret_cast<uint2*>(wide + (size_t)i * 32) =
            make_uint2(lo == gamma ? (0x80000000u | gamma) : gamma, hi == gamma + 1 ? (0x80000000u | (gamma + 1)) : gamma + 1);
        // arrival word of the refit (refit.cu): marked when all leaves of the node lie in one
        // 256-leaf window, i.e. inside one refit block, which then resolves it in shared memory
        flags[i] = ((lo >> 8) == (hi >> 8)) ? 0x80000000u : 0u;
    }
}

__global__ void __launch_bounds__(256)
fill_leaves_kernel(col_node* __restrict__ nodes, const uint32_t* __restrict__ ids, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    nodes[n - 1 + i].data[0] = ids[i];
    nodes[n - 1 + i].right_edge = i;
}

}  // namespace

COL_API int col_build_tree(const uint32_t* codes_sorted, const uint32_t* ids_sorted, void* nodes, uint32_t n,
                           void* stream) {
    if (!codes_sorted || !nodes) return COL_ERR_BAD_ARG;
    if (n == 0) return 0;
    if (n > 0x80000000u) return COL_ERR_TOO_LARGE;
    const unsigned blocks = col_blocks(n, 256);
    if (ids_sorted)
        build_tree_kernel<true, false><<<blocks, 256, 0, col_stream(stream)>>>(codes_sorted, ids_sorted, (col_node*)nodes, n, nullptr, nullptr);
    else
        build_tree_kernel<false, false><<<blocks, 256, 0, col_stream(stream)>>>(codes_sorted, nullptr, (col_node*)nodes, n, nullptr, nullptr);
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_build_tree_wide(const uint32_t* codes_sorted, const uint32_t* ids_sorted, void* nodes, void* wide,
                                uint32_t* flags, uint32_t n, void* stream) {
    if (!codes_sorted || !ids_sorted || !nodes || !wide || !flags) return COL_ERR_BAD_ARG;
    if (n == 0) return 0;
    if (n > 0x80000000u) return COL_ERR_TOO_LARGE;
    build_tree_kernel<true, true><<<col_blocks(n, 256), 256, 0, col_stream(stream)>>>(
        codes_sorted, ids_sorted, (col_node*)nodes, n, (unsigned char*)wide, flags);
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_fill_leaves(void* nodes, const uint32_t* ids_sorted, uint32_t n, void* stream) {
    if (!nodes || !ids_sorted) return COL_ERR_BAD_ARG;
    if (n == 0) return 0;
    fill_leaves_kernel<<<col_blocks(n, 256), 256, 0, col_stream(stream)>>>((col_node*)nodes, ids_sorted, n);
    COL_CHECK_LAUNCH();
    return 0;
}
