// refit.cu -- leaf boxes + bottom-up AABB refit with per-node arrival flags.
//
// Replaces leafBounds + internalBounds (collision/collision.cl:128-162, two
// launches upstream) with one: thread i computes the box of sorted leaf i
// (coords[id] -/+ radii[id], IEEE add/sub), stores it, then walks towards the
// root.  At each internal node an atomic arrival counter decides: the first
// thread to arrive stops, the second merges the two child boxes (component
// min / max -- exact and order-free, hence bit-identical to upstream) and
// continues.  The surviving thread carries its own child's box in registers
// and only reads the sibling's from memory (L2-coherent ld.global.cg after a
// __threadfence on both sides), halving the box re-reads of the reference.
#include "common.cuh"
#include "qnode.cuh"

namespace {

template <typename T> struct Box { T lo[3], hi[3]; };

__device__ __forceinline__ void store_box(float* bounds, size_t node, const Box<float>& b) {
    float4* p = reinterpret_cast<float4*>(bounds + node * 8);
    p[0] = make_float4(b.lo[0], b.lo[1], b.lo[2], 0.0f);
    p[1] = make_float4(b.hi[0], b.hi[1], b.hi[2], 0.0f);
}
__device__ __forceinline__ void store_box(double* bounds, size_t node, const Box<double>& b) {
    double2* p = reinterpret_cast<double2*>(bounds + node * 8);
    p[0] = make_double2(b.lo[0], b.lo[1]);
    p[1] = make_double2(b.lo[2], 0.0);
    p[2] = make_double2(b.hi[0], b.hi[1]);
    p[3] = make_double2(b.hi[2], 0.0);
}
__device__ __forceinline__ Box<float> load_box_cg(const float* bounds, size_t node) {
    const float4* p = reinterpret_cast<const float4*>(bounds + node * 8);
    const float4 a = ld_cg_f4(p), c = ld_cg_f4(p + 1);
    Box<float> b;
    b.lo[0] = a.x; b.lo[1] = a.y; b.lo[2] = a.z;
    b.hi[0] = c.x; b.hi[1] = c.y; b.hi[2] = c.z;
    return b;
}
__device__ __forceinline__ Box<double> load_box_cg(const double* bounds, size_t node) {
    const double2* p = reinterpret_cast<const double2*>(bounds + node * 8);
    const double2 a = ld_cg_d2(p), c = ld_cg_d2(p + 1), d = ld_cg_d2(p + 2), e = ld_cg_d2(p + 3);
    Box<double> b;
    b.lo[0] = a.x; b.lo[1] = a.y; b.lo[2] = c.x;
    b.hi[0] = d.x; b.hi[1] = d.y; b.hi[2] = e.x;
    return b;
}
__device__ __forceinline__ float vmin(float a, float b) { return fminf(a, b); }
__device__ __forceinline__ float vmax(float a, float b) { return fmaxf(a, b); }
__device__ __forceinline__ double vmin(double a, double b) { return fmin(a, b); }
__device__ __forceinline__ double vmax(double a, double b) { return fmax(a, b); }

// one arrival on a node's counter with release + acquire semantics at GPU scope (cheaper than
// the two sequentially-consistent __threadfence() around a relaxed atomic it replaces)
__device__ __forceinline__ uint32_t arrive_acq_rel(uint32_t* p) {
    uint32_t old;
    asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(p) : "memory");
    return old;
}

__device__ __forceinline__ uint32_t arrive_acq_rel_shared(unsigned* p) {
    uint32_t old;
    asm volatile("atom.add.acq_rel.cta.shared.u32 %0, [%1], 1;"
                 : "=r"(old)
                 : "r"((uint32_t)__cvta_generic_to_shared(p))
                 : "memory");
    return old;
}

template <typename T>
__device__ __forceinline__ Box<T> leaf_box(const T* __restrict__ coords, const T* __restrict__ radii, uint32_t id) {
    Box<T> b;
    T c[3];
    if constexpr (sizeof(T) == 4) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(coords) + id);
        c[0] = v.x; c[1] = v.y; c[2] = v.z;
    } else {
        const double2* p = reinterpret_cast<const double2*>(coords) + (size_t)id * 2;
        const double2 a = __ldg(p), d = __ldg(p + 1);
        c[0] = a.x; c[1] = a.y; c[2] = d.x;
    }
    const T r = __ldg(radii + id);
#pragma unroll
    for (int k = 0; k < 3; k++) {
        b.lo[k] = c[k] - r;
        b.hi[k] = c[k] + r;
    }
    return b;
}

// LEAF: compute + store the leaf boxes first (coords/radii given); otherwise the
// leaf boxes are already in `bounds` (kernel-level internalBounds contract).
template <typename T, bool LEAF>
__global__ void __launch_bounds__(256)
refit_kernel(T* __restrict__ bounds, uint32_t* __restrict__ flags, const col_node* __restrict__ nodes,
             const T* __restrict__ coords, const T* __restrict__ radii, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t leaf_start = n - 1;
    uint32_t node = leaf_start + i;
    // one 16-byte load: {parent, right_edge, id, -}
    uint4 rec = __ldg(reinterpret_cast<const uint4*>(nodes + node));
    Box<T> box;
    if (LEAF) {
        box = leaf_box<T>(coords, radii, rec.z);
        store_box(bounds, node, box);
    } else {
        box = load_box_cg(bounds, node);
    }
    if (n < 2) return;
    while (true) {
        const uint32_t parent = rec.x;
        // my child's box is visible before my arrival is (release), the sibling's after it (acquire)
        if (arrive_acq_rel(flags + parent) < 1u) return;  // first to arrive: the sibling will carry on
        rec = __ldg(reinterpret_cast<const uint4*>(nodes + parent));  // {grandparent, right_edge, child a, child b}
        const uint32_t sibling = rec.z == node ? rec.w : rec.z;
        const Box<T> other = load_box_cg(bounds, sibling);
#pragma unroll
        for (int k = 0; k < 3; k++) {
            box.lo[k] = vmin(box.lo[k], other.lo[k]);
            box.hi[k] = vmax(box.hi[k], other.hi[k]);
        }
        store_box(bounds, parent, box);
        node = parent;
        if (node == 0) return;
    }
}

// ---- refit emitting the traversal records (qnode.cuh) -------------------------
// One block = 256 consecutive sorted leaves.  An internal node all of whose leaves lie
// in that window (marked by the tree kernel in flags[], 255 of every 256 nodes) is
// resolved entirely inside the block: both arrivals leave their exact box in shared
// memory (block-scope fence + shared-memory counter), the second merges, quantises BOTH
// child boxes and writes the node's traversal record in one go -- no global atomic, no
// device-scope fence.  Only nodes spanning several windows take the global path, where
// the sibling's exact box is read back from the Bound array.
constexpr uint32_t REFIT_IN_BLOCK = 0x80000000u;

template <typename T>
__global__ void __launch_bounds__(256)
refit_wide_kernel(T* __restrict__ bounds, unsigned char* __restrict__ wide, uint32_t* __restrict__ flags,
                  const col_node* __restrict__ nodes, const T* __restrict__ coords, const T* __restrict__ radii,
                  const T* __restrict__ scene, uint32_t n) {
    __shared__ T s_box[256][2][6];   // [node & 255][which child][lo xyz, hi xyz]
    __shared__ unsigned s_cnt[256];
    s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Quant<T> qz = make_quant(scene);
    const uint32_t leaf_start = n - 1;
    uint32_t node = leaf_start + i;
    const uint4 leaf = __ldg(reinterpret_cast<const uint4*>(nodes + node));  // {parent, right_edge, id, -}
    Box<T> box = leaf_box<T>(coords, radii, leaf.z);
    store_box(bounds, node, box);
    if (n < 2) return;
    uint32_t parent = leaf.x;
    while (true) {
        const uint4 rec = __ldg(reinterpret_cast<const uint4*>(nodes + parent));  // {grandparent, right_edge, a, b}
        const int half = rec.z == node ? 0 : 1;
        Box<T> other;
        if (__ldg(flags + parent) & REFIT_IN_BLOCK) {
            const uint32_t slot = parent & 255u;
#pragma unroll
            for (int k = 0; k < 3; k++) {
                s_box[slot][half][k] = box.lo[k];
                s_box[slot][half][3 + k] = box.hi[k];
            }
            if (arrive_acq_rel_shared(&s_cnt[slot]) == 0u) return;  // first to arrive: the sibling carries on
#pragma unroll
            for (int k = 0; k < 3; k++) {
                other.lo[k] = s_box[slot][half ^ 1][k];
                other.hi[k] = s_box[slot][half ^ 1][3 + k];
            }
        } else {
            // my exact box is already in the Bound array (stored below / at the leaf)
            // ... and visible before my arrival is (release); the sibling's box after it (acquire)
            if (arrive_acq_rel(flags + parent) < 1u) return;
            other = load_box_cg(bounds, half == 0 ? rec.w : rec.z);
        }
        {   // the node's traversal record: both children, quantised outward
            const uint3 mine = quantise_box(qz, box.lo, box.hi), sib = quantise_box(qz, other.lo, other.hi);
            const uint3 qa = half == 0 ? mine : sib, qb = half == 0 ? sib : mine;
            unsigned char* rp = wide + (size_t)parent * QNODE_BYTES;
            *reinterpret_cast<uint2*>(rp + 8) = make_uint2(qa.x, qa.y);
            *reinterpret_cast<uint4*>(rp + 16) = make_uint4(qa.z, qb.x, qb.y, qb.z);
        }
#pragma unroll
        for (int k = 0; k < 3; k++) {
            box.lo[k] = vmin(box.lo[k], other.lo[k]);
            box.hi[k] = vmax(box.hi[k], other.hi[k]);
        }
        store_box(bounds, parent, box);
        node = parent;
        if (node == 0) return;
        parent = rec.x;
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
leaf_bounds_kernel(T* __restrict__ bounds, const T* __restrict__ coords, const T* __restrict__ radii,
                   const col_node* __restrict__ nodes, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t node = (size_t)(n - 1) + i;
    store_box(bounds, node, leaf_box<T>(coords, radii, nodes[node].data[0]));
}

}  // namespace

COL_API int col_refit(void* bounds, uint32_t* flags, const void* nodes, const void* coords4, const void* radii,
                      int dtype, uint32_t n, void* stream) {
    if (!bounds || !flags || !nodes) return COL_ERR_BAD_ARG;
    if ((coords4 == nullptr) != (radii == nullptr)) return COL_ERR_BAD_ARG;
    if (n == 0) return 0;
    cudaStream_t s = col_stream(stream);
    COL_CUDA(cudaMemsetAsync(flags, 0, sizeof(uint32_t) * (2 * (size_t)n - 1), s));
    const unsigned blocks = col_blocks(n, 256);
    const bool leaf = coords4 != nullptr;
    if (dtype == COL_F32) {
        if (leaf) refit_kernel<float, true><<<blocks, 256, 0, s>>>((float*)bounds, flags, (const col_node*)nodes, (const float*)coords4, (const float*)radii, n);
        else refit_kernel<float, false><<<blocks, 256, 0, s>>>((float*)bounds, flags, (const col_node*)nodes, nullptr, nullptr, n);
    } else if (dtype == COL_F64) {
        if (leaf) refit_kernel<double, true><<<blocks, 256, 0, s>>>((double*)bounds, flags, (const col_node*)nodes, (const double*)coords4, (const double*)radii, n);
        else refit_kernel<double, false><<<blocks, 256, 0, s>>>((double*)bounds, flags, (const col_node*)nodes, nullptr, nullptr, n);
    } else {
        return COL_ERR_UNSUPPORTED;
    }
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_refit_wide(void* bounds, void* wide, uint32_t* flags, const void* nodes, const void* coords4,
                           const void* radii, const void* scene2x4, int dtype, uint32_t n, void* stream) {
    if (!bounds || !wide || !flags || !nodes || !coords4 || !radii || !scene2x4) return COL_ERR_BAD_ARG;
    if (scene2x4 == bounds) return COL_ERR_BAD_ARG;  // slot 0 of the Bound array is overwritten with the root box
    if (n == 0) return 0;
    cudaStream_t s = col_stream(stream);
    // flags[] were initialised by col_build_tree_wide (0 or the in-block mark): no memset here
    const unsigned blocks = col_blocks(n, 256);
    if (dtype == COL_F32)
        refit_wide_kernel<float><<<blocks, 256, 0, s>>>((float*)bounds, (unsigned char*)wide, flags, (const col_node*)nodes, (const float*)coords4, (const float*)radii, (const float*)scene2x4, n);
    else if (dtype == COL_F64)
        refit_wide_kernel<double><<<blocks, 256, 0, s>>>((double*)bounds, (unsigned char*)wide, flags, (const col_node*)nodes, (const double*)coords4, (const double*)radii, (const double*)scene2x4, n);
    else
        return COL_ERR_UNSUPPORTED;
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_leaf_bounds(void* bounds, const void* coords4, const void* radii, const void* nodes, int dtype,
                            uint32_t n, void* stream) {
    if (!bounds || !coords4 || !radii || !nodes) return COL_ERR_BAD_ARG;
    if (n == 0) return 0;
    const unsigned blocks = col_blocks(n, 256);
    if (dtype == COL_F32)
        leaf_bounds_kernel<float><<<blocks, 256, 0, col_stream(stream)>>>((float*)bounds, (const float*)coords4, (const float*)radii, (const col_node*)nodes, n);
    else if (dtype == COL_F64)
        leaf_bounds_kernel<double><<<blocks, 256, 0, col_stream(stream)>>>((double*)bounds, (const double*)coords4, (const double*)radii, (const col_node*)nodes, n);
    else
        return COL_ERR_UNSUPPORTED;
    COL_CHECK_LAUNCH();
    return 0;
}
