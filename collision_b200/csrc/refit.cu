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

// ---- tree + refit in one bottom-up pass ----------------------------------------
// The radix tree over the (code, position) keys is unique, so it can be found from the
// leaves upwards (Apetrei 2014) instead of by Karras' per-node searches, in the same climb
// that merges the boxes.  A node covering sorted positions [lo, hi] is the LEFT child of its
// parent iff its right neighbour shares the longer prefix, delta(hi, hi+1) > delta(lo-1, lo)
// (delta as collision.cl:55-70, -1 outside the array; the two are never equal inside a sorted
// array of distinct keys); the parent's split then lies at hi, else at lo-1.  Both children of
// a node meet on the word flags[split]: the first to arrive leaves the far end of its range
// there and stops, the second picks it up and so knows the parent's range.
// Karras' numbering follows from the ranges alone: the left child of a split at g is node g
// (its hi), the right child node g+1 (its lo), the root node 0 -- so every field of the
// reference's Node array (collision.cl:72-121) is reproduced: .right_edge = hi, .data = the two
// children, .parent written by the parent's thread.
constexpr uint32_t NO_ARRIVAL = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t exchange_acq_rel(uint32_t* p, uint32_t v) {
    uint32_t old;
    asm volatile("atom.exch.acq_rel.gpu.global.b32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ uint32_t exchange_acq_rel_shared(uint32_t* p, uint32_t v) {
    uint32_t old;
    asm volatile("atom.exch.acq_rel.cta.shared.b32 %0, [%1], %2;"
                 : "=r"(old)
                 : "r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v)
                 : "memory");
    return old;
}
// delta of sorted positions a < b (collision.cl:55-70 for in-range arguments)
__device__ __forceinline__ int key_delta(uint32_t a, uint32_t code_a, uint32_t b, uint32_t code_b) {
    const uint32_t x = code_a ^ code_b;
    return x ? __clz(x) : 32 + __clz(a ^ b);
}

// One block = 256 consecutive sorted leaves [first, last].  A parent all of whose leaves lie in
// that window is resolved inside the block (boxes and arrival word in shared memory, no global
// atomic); whether it does is known from the child alone: the parent is the set of keys
// sharing its prefix, of length delta(split, split+1), with the child, so it reaches past the
// window's end iff delta(hi, last+1) is at least that long (likewise first-1 on the left).
constexpr uint32_t BR_WINDOW = 256;   // leaves (= threads) per block of the window pass
#ifndef COL_BR_TWO_PASS_MIN_N
#define COL_BR_TWO_PASS_MIN_N (1u << 22)
#endif

// What the second arrival at a node does once it knows the node: [lo, hi] (already extended by
// the sibling's range), both children and their boxes.  Writes the node's record of the Node
// array, both children's .parent, its traversal record and its exact box; returns its index.
template <typename T>
__device__ __forceinline__ uint32_t finish_node(col_node* __restrict__ nodes, unsigned char* __restrict__ wide,
                                                T* __restrict__ bounds, uint32_t leaf_start, uint32_t lo, uint32_t hi,
                                                int dl, int dr, bool is_left, uint32_t self, uint32_t sibling,
                                                Box<T>& box, const Box<T>& other, uint3& grid, const uint3& other_grid) {
    const bool root = lo == 0 && hi == leaf_start;
    const uint32_t idx = root ? 0u : (dr > dl ? hi : lo);  // Karras: a left child is numbered by its hi, a right child by its lo
    const uint32_t child_a = is_left ? self : sibling, child_b = is_left ? sibling : self;
#ifdef COL_DEBUG
    if (lo > hi || hi > leaf_start || idx >= leaf_start || child_a > 2 * leaf_start || child_b > 2 * leaf_start) __trap();
#endif
    {
        uint32_t* rec = reinterpret_cast<uint32_t*>(nodes + idx);
        rec[1] = hi;
        *reinterpret_cast<uint2*>(rec + 2) = make_uint2(child_a, child_b);
        nodes[child_a].parent = idx;
        nodes[child_b].parent = idx;
    }
    {   // the node's traversal record: children (a leaf as 0x80000000 | position) + their grid boxes
        const uint3 qa = is_left ? grid : other_grid, qb = is_left ? other_grid : grid;
        const uint32_t wa = child_a >= leaf_start ? (0x80000000u | (child_a - leaf_start)) : child_a;
        const uint32_t wb = child_b >= leaf_start ? (0x80000000u | (child_b - leaf_start)) : child_b;
        unsigned char* rp = wide + (size_t)idx * QNODE_BYTES;
        *reinterpret_cast<uint4*>(rp) = make_uint4(wa, wb, qa.x, qa.y);
        *reinterpret_cast<uint4*>(rp + 16) = make_uint4(qa.z, qb.x, qb.y, qb.z);
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
        box.lo[k] = vmin(box.lo[k], other.lo[k]);
        box.hi[k] = vmax(box.hi[k], other.hi[k]);
    }
    store_box(bounds, idx, box);
    grid = grid_union(grid, other_grid);
    return idx;
}

// ---- the nodes that span windows: arrival words in `flags` (device-scope exchange), the
// sibling's exact box read back from the Bound array.  Climbs from the subtree [lo, hi] = node
// `self` until it is the first to arrive somewhere, or at the root.
template <typename T>
__device__ __forceinline__ void top_climb(const uint32_t* __restrict__ codes, col_node* __restrict__ nodes,
                                          unsigned char* __restrict__ wide, uint32_t* __restrict__ flags,
                                          T* __restrict__ bounds, const Quant<T>& qz, uint32_t leaf_start, uint32_t lo,
                                          uint32_t hi, uint32_t self, uint32_t code_lo, uint32_t code_hi, int dl, int dr,
                                          Box<T> box, uint3 grid) {
    while (true) {
        const bool is_left = dr > dl;
        const uint32_t split = is_left ? hi : lo - 1;
        // my box is in the Bound array before my arrival is visible (release); the sibling's after it (acquire)
        const uint32_t far_end = exchange_acq_rel(flags + split, is_left ? lo : hi);
        if (far_end == NO_ARRIVAL) return;  // first to arrive: the sibling carries on
        uint32_t sibling;
        if (is_left) {
            sibling = far_end == split + 1 ? leaf_start + split + 1 : split + 1;
            hi = far_end;
            code_hi = __ldg(codes + hi);
            dr = hi < leaf_start ? key_delta(hi, code_hi, hi + 1, __ldg(codes + hi + 1)) : -1;
        } else {
            sibling = far_end == split ? leaf_start + split : split;
            lo = far_end;
            code_lo = __ldg(codes + lo);
            dl = lo > 0 ? key_delta(lo - 1, __ldg(codes + lo - 1), lo, code_lo) : -1;
        }
        const Box<T> other = load_box_cg(bounds, sibling);
        const uint3 other_grid = quantise_box(qz, other.lo, other.hi);
        self = finish_node(nodes, wide, bounds, leaf_start, lo, hi, dl, dr, is_left, self, sibling, box, other, grid, other_grid);
        if (lo == 0 && hi == leaf_start) return;
    }
}

// ---- the climb from the leaves: everything inside the 256-leaf windows through shared memory,
// the rest through `flags` and the Bound array (same thread, or handed over)
// HAND_OVER (large scenes): a subtree root whose parent reaches beyond the window is handed to a
// second launch through the block's slots of `items`, so that the block's registers and shared
// memory are not held while the few cross-window climbs (6 % of the arrivals, each a chain of
// device-scope round trips) run their course.  Otherwise (small scenes) the climb simply
// continues in the same thread: its latency then overlaps the other blocks' window work.
template <typename T, bool HAND_OVER>
__global__ void __launch_bounds__(256, 5)
build_refit_kernel(const uint32_t* __restrict__ codes, const uint32_t* __restrict__ ids, col_node* __restrict__ nodes,
                   unsigned char* __restrict__ wide, uint32_t* __restrict__ flags, uint32_t* __restrict__ items,
                   T* __restrict__ bounds,
                   const T* __restrict__ coords, const T* __restrict__ radii, const T* __restrict__ scene,
                   uint32_t n, uint32_t first_window) {
    __shared__ T s_box[256][2][6];          // [split & 255][which child][lo xyz, hi xyz]
    __shared__ uint32_t s_grid[256][2][3];  // ... and its grid box
    __shared__ uint32_t s_far[256];         // arrival word of the in-window splits
    __shared__ uint32_t s_code[258];        // codes of positions first-1 .. last+1
    __shared__ uint32_t s_items;            // subtree roots handed on so far (HAND_OVER)
    const uint32_t first = (first_window + blockIdx.x) * blockDim.x;   // (first_window: the sharded build runs a rank's own windows only)
    const uint32_t i = first + threadIdx.x;
    s_far[threadIdx.x] = NO_ARRIVAL;
    if (threadIdx.x == 0) s_items = 0;
    if (i < n) s_code[threadIdx.x + 1] = __ldg(codes + i);
    if (threadIdx.x == 0) s_code[0] = first > 0 ? __ldg(codes + first - 1) : 0u;
    if (threadIdx.x == 32) s_code[257] = first + 256 < n ? __ldg(codes + first + 256) : 0u;
    __syncthreads();
    if (i >= n) return;
    const Quant<T> qz = make_quant(scene);
    const uint32_t leaf_start = n - 1;
    const uint32_t last = min(first + 255u, leaf_start);
    const uint32_t id = __ldg(ids + i);
    Box<T> box = leaf_box<T>(coords, radii, id);
    uint32_t self = leaf_start + i;
    store_box(bounds, self, box);
    // grid box of the current node: quantised once here (all lanes busy), merged with packed minima on the way up
    uint3 grid = quantise_box(qz, box.lo, box.hi);
    {   // leaf record {parent (by the parent's thread), right_edge, id, NO_NODE}
        uint32_t* rec = reinterpret_cast<uint32_t*>(nodes + self);
        rec[1] = i;
        rec[2] = id;
        rec[3] = 0xFFFFFFFFu;
    }
    if (i == 0) nodes[0].parent = 0xFFFFFFFFu;
    if (n < 2) return;
    uint32_t lo = i, hi = i;
    uint32_t code_lo = s_code[threadIdx.x + 1], code_hi = code_lo;
    // prefix shared with the left / right neighbour of the range (-1: none)
    int dl = lo > 0 ? key_delta(lo - 1, s_code[threadIdx.x], lo, code_lo) : -1;
    int dr = hi < leaf_start ? key_delta(hi, code_hi, hi + 1, s_code[threadIdx.x + 2]) : -1;
    bool local = true;  // the current node lies inside the block's window
    while (true) {
        const bool is_left = dr > dl;
        const uint32_t split = is_left ? hi : lo - 1;
        if (local) {
            if (is_left) local = last == leaf_start || key_delta(hi, code_hi, last + 1, s_code[last - first + 2]) < dr;
            else local = first == 0 || key_delta(first - 1, s_code[0], lo, code_lo) < dl;
        }
        Box<T> other;
        uint3 other_grid;
        uint32_t far_end, end_code, next_code;  // the sibling's far end, its code, the code one position beyond
        if (local) {
            const uint32_t slot = split & 255u;
            const int half = is_left ? 0 : 1;
#pragma unroll
            for (int k = 0; k < 3; k++) {
                s_box[slot][half][k] = box.lo[k];
                s_box[slot][half][3 + k] = box.hi[k];
            }
            s_grid[slot][half][0] = grid.x;
            s_grid[slot][half][1] = grid.y;
            s_grid[slot][half][2] = grid.z;
            far_end = exchange_acq_rel_shared(&s_far[slot], is_left ? lo : hi);
            if (far_end == NO_ARRIVAL) return;  // first to arrive: the sibling carries on
#pragma unroll
            for (int k = 0; k < 3; k++) {
                other.lo[k] = s_box[slot][half ^ 1][k];
                other.hi[k] = s_box[slot][half ^ 1][3 + k];
            }
            other_grid = make_uint3(s_grid[slot][half ^ 1][0], s_grid[slot][half ^ 1][1], s_grid[slot][half ^ 1][2]);
            // all leaves of an in-window parent, and their neighbours, are among the staged codes
            const uint32_t k = far_end - first + 1;
            end_code = s_code[k];
            next_code = s_code[is_left ? k + 1 : k - 1];
        } else {
            if (HAND_OVER) {
                // the subtree's node index; an internal node's lo travels in its own .parent field,
                // which nobody writes before both children of its parent have arrived
                if (self < leaf_start) nodes[self].parent = lo;
                items[first + atomicAdd(&s_items, 1u)] = self;
                return;
            }
            // my box is in the Bound array before my arrival is visible (release); the sibling's after it (acquire)
            far_end = exchange_acq_rel(flags + split, is_left ? lo : hi);
            if (far_end == NO_ARRIVAL) return;
            const uint32_t sib = is_left ? (far_end == split + 1 ? leaf_start + split + 1 : split + 1)
                                         : (far_end == split ? leaf_start + split : split);
            end_code = __ldg(codes + far_end);
            next_code = is_left ? (far_end < leaf_start ? __ldg(codes + far_end + 1) : 0u)
                                : (far_end > 0 ? __ldg(codes + far_end - 1) : 0u);
            other = load_box_cg(bounds, sib);
            other_grid = quantise_box(qz, other.lo, other.hi);
        }
        uint32_t sibling;
        if (is_left) {
            sibling = far_end == split + 1 ? leaf_start + split + 1 : split + 1;
            hi = far_end;
            code_hi = end_code;
            dr = hi < leaf_start ? key_delta(hi, code_hi, hi + 1, next_code) : -1;
        } else {
            sibling = far_end == split ? leaf_start + split : split;
            lo = far_end;
            code_lo = end_code;
            dl = lo > 0 ? key_delta(lo - 1, next_code, lo, code_lo) : -1;
        }
        const bool root = lo == 0 && hi == leaf_start;
        const uint32_t idx = root ? 0u : (dr > dl ? hi : lo);  // Karras: a left child is numbered by its hi, a right child by its lo
        const uint32_t child_a = is_left ? self : sibling, child_b = is_left ? sibling : self;
#ifdef COL_DEBUG   // a torn hand-over (far end read before it was written) shows up as an impossible range or index
        if (lo > hi || hi > leaf_start || idx >= leaf_start || child_a > 2 * leaf_start || child_b > 2 * leaf_start) __trap();
#endif
        {
            uint32_t* rec = reinterpret_cast<uint32_t*>(nodes + idx);
            rec[1] = hi;
            *reinterpret_cast<uint2*>(rec + 2) = make_uint2(child_a, child_b);
            nodes[child_a].parent = idx;
            nodes[child_b].parent = idx;
        }
        {   // the node's traversal record: children (a leaf as 0x80000000 | position) + their grid boxes
            const uint3 qa = is_left ? grid : other_grid, qb = is_left ? other_grid : grid;
            const uint32_t wa = child_a >= leaf_start ? (0x80000000u | (child_a - leaf_start)) : child_a;
            const uint32_t wb = child_b >= leaf_start ? (0x80000000u | (child_b - leaf_start)) : child_b;
            unsigned char* rp = wide + (size_t)idx * QNODE_BYTES;
            *reinterpret_cast<uint4*>(rp) = make_uint4(wa, wb, qa.x, qa.y);
            *reinterpret_cast<uint4*>(rp + 16) = make_uint4(qa.z, qb.x, qb.y, qb.z);
        }
#pragma unroll
        for (int k = 0; k < 3; k++) {
            box.lo[k] = vmin(box.lo[k], other.lo[k]);
            box.hi[k] = vmax(box.hi[k], other.hi[k]);
        }
        store_box(bounds, idx, box);
        if (root) return;
        grid = grid_union(grid, other_grid);
        self = idx;
    }
}

// ---- pass 2 (large scenes): one thread per hand-over slot of pass 1
template <typename T>
__global__ void __launch_bounds__(256)
build_refit_top_kernel(const uint32_t* __restrict__ codes, col_node* __restrict__ nodes,
                       unsigned char* __restrict__ wide, uint32_t* __restrict__ flags,
                       const uint32_t* __restrict__ items, T* __restrict__ bounds, const T* __restrict__ scene,
                       uint32_t n) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const uint32_t self = items[t];
    if (self == NO_ARRIVAL) return;   // slot not used
    const Quant<T> qz = make_quant(scene);
    const uint32_t leaf_start = n - 1;
    uint32_t lo, hi;
    if (self >= leaf_start) {
        lo = hi = self - leaf_start;
    } else {
        const uint2 rec = *reinterpret_cast<const uint2*>(nodes + self);   // {lo (see pass 1), right_edge}
        lo = rec.x;
        hi = rec.y;
    }
    const Box<T> box = load_box_cg(bounds, self);
    const uint32_t code_lo = __ldg(codes + lo), code_hi = __ldg(codes + hi);
    const int dl = lo > 0 ? key_delta(lo - 1, __ldg(codes + lo - 1), lo, code_lo) : -1;
    const int dr = hi < leaf_start ? key_delta(hi, code_hi, hi + 1, __ldg(codes + hi + 1)) : -1;
    top_climb(codes, nodes, wide, flags, bounds, qz, leaf_start, lo, hi, self, code_lo, code_hi, dl, dr, box,
              quantise_box(qz, box.lo, box.hi));
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


// ================================================================ sharded build (multi-GPU)
// One process per GPU (collision_b200/sharded.py, mode "sharded_build").  The windows of the
// sorted leaves are dealt out to the ranks; a rank runs the window pass for ITS windows only, the
// roots it hands over are packed and all-gathered (a few per window), every rank finishes the few
// nodes that span windows itself (top pass, replicated: it is small), and the in-window records
// and leaf boxes of OTHER ranks' windows are fetched on demand -- only the subtrees that the
// rank's own queries can enter -- straight from the owner's HBM over NVLink (peer loads; the
// owners' buffers are mapped through CUDA IPC, ipc.cu).
template <typename T> struct PackedItem {
    uint32_t self, lo, hi, pad;   // node index (leaf: leaf_start + position), covered sorted positions [lo, hi]
    uint32_t code[4];             // sorted codes at lo - 1, lo, hi, hi + 1 (0 where there is no such position): with the
                                  // sharded sort the other ranks do not hold them, and the top pass reads nothing else
    T box[6];                     // exact box: lo xyz, hi xyz
};
constexpr uint32_t PACK_HEADER = 64;   // bytes: {count, overflow, ...} in front of a rank's packed items

// own hand-over slots -> dense list (order arbitrary: the top pass does not depend on it)
template <typename T>
__global__ void __launch_bounds__(256)
pack_items_kernel(const uint32_t* __restrict__ items, const col_node* __restrict__ nodes, const T* __restrict__ bounds,
                  const uint32_t* __restrict__ codes, unsigned char* __restrict__ packed, uint32_t cap, uint32_t n,
                  uint32_t leaf_begin, uint32_t leaf_end) {
    const uint32_t s = leaf_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= leaf_end) return;
    const uint32_t self = items[s];
    if (self == NO_ARRIVAL) return;
    uint32_t* hdr = reinterpret_cast<uint32_t*>(packed);
    const uint32_t k = atomicAdd(hdr, 1u);
    if (k >= cap) {
        hdr[1] = 1u;   // overflow: reported by the host at its next synchronisation
        return;
    }
    const uint32_t leaf_start = n - 1;
    PackedItem<T> it;
    it.self = self;
    if (self >= leaf_start) {
        it.lo = it.hi = self - leaf_start;
    } else {
        const uint2 rec = *reinterpret_cast<const uint2*>(nodes + self);   // {lo (window pass), right_edge}
        it.lo = rec.x;
        it.hi = rec.y;
    }
    it.pad = 0;
    it.code[0] = it.lo > 0 ? __ldg(codes + it.lo - 1) : 0u;
    it.code[1] = __ldg(codes + it.lo);
    it.code[2] = __ldg(codes + it.hi);
    it.code[3] = it.hi < leaf_start ? __ldg(codes + it.hi + 1) : 0u;
    const T* b = bounds + (size_t)self * 8;
    it.box[0] = b[0]; it.box[1] = b[1]; it.box[2] = b[2];
    it.box[3] = b[4]; it.box[4] = b[5]; it.box[5] = b[6];
    reinterpret_cast<PackedItem<T>*>(packed + PACK_HEADER)[k] = it;
}

// Did any rank hand over more roots than the exchange holds?  Then the spanning nodes cannot be completed, and
// nothing may walk them (a walk over half-built records need not terminate): the top pass writes a root whose
// children are two copies of leaf 0 behind boxes that nothing but a query spanning the whole scene overlaps, the
// halo asks for nothing, the traversal ends at once, and the host -- which sees the same flags when it next
// syncs -- repeats the call with larger buffers or raises.
__device__ __forceinline__ bool shard_exchange_overflowed(const unsigned char* __restrict__ gathered, size_t stride,
                                                          uint32_t world) {
    uint32_t over = 0;
    for (uint32_t r = 0; r < world; r++) {
        const uint32_t* h = reinterpret_cast<const uint32_t*>(gathered + r * stride);
        over |= h[1] | h[3];   // {roots, roots overflow (or the rank's probe list), spanning nodes, their overflow}
    }
    return over != 0;
}

// What a rank puts into the all-gather: a 64-byte header {roots, overflow, spanning nodes, overflow}, up to XCHG_ROOTS
// roots of its maximal subtrees that stay inside its own range of leaves (PackedItem), and the indices of the nodes
// it built ABOVE its windows (they span windows, but not ranks): the other ranks copy those records from its HBM.
constexpr uint32_t XCHG_HEADER = 64;
constexpr uint32_t XCHG_ROOTS = 2048;   // (a range has at most two such subtrees per level of the tree)
template <typename T> __host__ __device__ constexpr size_t xchg_span_offset() {
    return XCHG_HEADER + (size_t)XCHG_ROOTS * sizeof(PackedItem<T>);
}

// The codes at the ends of the other ranks' roots -> their places in this rank's array of sorted codes
// (with the sharded sort it holds the rank's own part only).  Every position the top pass reads is the
// lo - 1 / lo / hi / hi + 1 of some root: a spanning node's range ends where two of its descendants' do.
template <typename T>
__global__ void __launch_bounds__(256)
scatter_root_codes_kernel(const unsigned char* __restrict__ gathered, size_t stride, uint32_t cap, uint32_t world,
                          uint32_t my_rank, uint32_t* __restrict__ codes, uint32_t n) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t r = (uint32_t)(t / cap), j = (uint32_t)(t % cap);
    if (r >= world || r == my_rank) return;
    if (shard_exchange_overflowed(gathered, stride, world)) return;
    const unsigned char* theirs = gathered + r * stride;
    if (j >= min(*reinterpret_cast<const uint32_t*>(theirs), cap)) return;
    const PackedItem<T>* it = reinterpret_cast<const PackedItem<T>*>(theirs + PACK_HEADER) + j;
    const uint32_t lo = it->lo, hi = it->hi;
    if (lo > 0) codes[lo - 1] = it->code[0];
    codes[lo] = it->code[1];
    codes[hi] = it->code[2];
    if (hi + 1 < n) codes[hi + 1] = it->code[3];
}

// top pass over the gathered roots of ALL ranks (replicated on every rank)
template <typename T>
__global__ void __launch_bounds__(256)
top_packed_kernel(const unsigned char* __restrict__ gathered, size_t stride, uint32_t cap, uint32_t world,
                  uint32_t my_rank, const uint32_t* __restrict__ codes, col_node* __restrict__ nodes,
                  unsigned char* __restrict__ wide, uint32_t* __restrict__ flags, T* __restrict__ bounds,
                  const T* __restrict__ scene, uint32_t n) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t r = (uint32_t)(t / cap), j = (uint32_t)(t % cap);
    if (r >= world) return;
    if (shard_exchange_overflowed(gathered, stride, world)) {
        if (t == 0) {   // (see shard_exchange_overflowed)
            uint4* rp = reinterpret_cast<uint4*>(wide);
            rp[0] = make_uint4(0x80000000u, 0x80000000u, 0xFFFFFFFFu, 0xFFFFFFFFu);
            rp[1] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
        }
        return;
    }
    const unsigned char* mine = gathered + r * stride;
    if (j >= min(*reinterpret_cast<const uint32_t*>(mine), cap)) return;
    const PackedItem<T> it = reinterpret_cast<const PackedItem<T>*>(mine + PACK_HEADER)[j];
    Box<T> box;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        box.lo[k] = it.box[k];
        box.hi[k] = it.box[3 + k];
    }
    // a sibling reads my box from the Bound array after my arrival: a remote root's box is not there yet
    if (r != my_rank) store_box(bounds, it.self, box);
    const Quant<T> qz = make_quant(scene);
    const uint32_t leaf_start = n - 1, lo = it.lo, hi = it.hi;
    const uint32_t code_lo = __ldg(codes + lo), code_hi = __ldg(codes + hi);
    const int dl = lo > 0 ? key_delta(lo - 1, __ldg(codes + lo - 1), lo, code_lo) : -1;
    const int dr = hi < leaf_start ? key_delta(hi, code_hi, hi + 1, __ldg(codes + hi + 1)) : -1;
    top_climb(codes, nodes, wide, flags, bounds, qz, leaf_start, lo, hi, it.self, code_lo, code_hi, dl, dr, box,
              quantise_box(qz, box.lo, box.hi));
}

// Which in-window subtrees of OTHER ranks can this rank's queries enter?  One thread per own root:
// its box contains the boxes of all the queries below it, so a walk of the nodes that span windows
// (all local after the top pass) with that box finds a superset of what those queries will visit.
// A child whose leaves lie in one window is where the walk stops: if the window is another rank's
// and nobody has asked for it yet, [first leaf, last leaf] goes on the fetch list.  Ranges follow
// from Karras' numbering alone: the children of a node covering [lo, hi] with split g cover
// [lo, g] and [g + 1, hi], g = the left child's index (internal) or position (leaf).
template <typename T>
__global__ void __launch_bounds__(256)
halo_probe_kernel(const unsigned char* __restrict__ gathered, size_t stride, uint32_t world,
                  const unsigned char* __restrict__ my_packed, uint32_t cap, const unsigned char* __restrict__ wide,
                  const T* __restrict__ scene, uint32_t* __restrict__ asked, uint2* __restrict__ list,
                  uint32_t* __restrict__ list_hdr, uint32_t list_cap, uint32_t n, uint32_t leaf_begin,
                  uint32_t leaf_end) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= min(*reinterpret_cast<const uint32_t*>(my_packed), cap)) return;
    if (shard_exchange_overflowed(gathered, stride, world)) return;
    const PackedItem<T> it = reinterpret_cast<const PackedItem<T>*>(my_packed + PACK_HEADER)[j];
    const Quant<T> qz = make_quant(scene);
    T blo[3] = {it.box[0], it.box[1], it.box[2]}, bhi[3] = {it.box[3], it.box[4], it.box[5]};
    const uint3 qg = quantise_query(qz, blo, bhi);
    const uint32_t leaf_start = n - 1, my_lo = it.lo;
    uint32_t st_idx[64], st_lo[64], st_hi[64];
    int sp = 0;
    uint32_t idx = 0, lo = 0, hi = leaf_start;
    if ((lo >> 8) == (hi >> 8)) return;   // a single window: nothing spans windows
#ifdef COL_DEBUG
    uint32_t visits = 0;
#endif
    while (true) {
#ifdef COL_DEBUG   // a walk over half-built spanning nodes need not end: trap instead of hanging
        if (++visits > 2u * n || idx >= leaf_start || lo > hi || hi > leaf_start) __trap();
#endif
        const uint4* rp = reinterpret_cast<const uint4*>(wide + (size_t)idx * QNODE_BYTES);
        const uint4 r0 = rp[0], r1 = rp[1];
        const bool la = (int32_t)r0.x < 0, lb = (int32_t)r0.y < 0;
        const uint32_t g = r0.x & 0x7fffffffu;   // the split: left child's index = its last leaf (or the leaf itself)
        // child a covers [lo, g], child b covers [g + 1, hi]; only leaves right of my first query matter
        bool oa = grid_overlap(qg, r0.z, r0.w, r1.x) && g > my_lo;
        bool ob = grid_overlap(qg, r1.y, r1.z, r1.w) && hi > my_lo;
        const bool wa = (lo >> 8) == (g >> 8), wb = ((g + 1) >> 8) == (hi >> 8);   // child inside one window
#pragma unroll
        for (int c = 0; c < 2; c++) {
            const bool o = c ? ob : oa, w = c ? wb : wa;
            const uint32_t cl = c ? g + 1 : lo, ch = c ? hi : g;
            if (o && w && (cl < leaf_begin || cl >= leaf_end)) {
                if (atomicCAS(asked + cl, NO_ARRIVAL, NO_ARRIVAL - 1u) == NO_ARRIVAL) {
                    const uint32_t k = atomicAdd(list_hdr, 1u);
                    // (bit 31 of .y: the subtree is a LEFT child -- its root is node ch, else node cl; see halo_fetch_kernel)
                    if (k < list_cap) list[k] = make_uint2(cl, ch | (c ? 0u : 0x80000000u));
                    else list_hdr[1] = 1u;
                }
            }
        }
        (void)la; (void)lb;
        const bool ta = oa && !wa, tb = ob && !wb;
        if (ta && tb) {
            st_idx[sp] = r0.y; st_lo[sp] = g + 1; st_hi[sp] = hi;
            sp++;
        }
        if (ta) { idx = r0.x; hi = g; }
        else if (tb) { idx = r0.y; lo = g + 1; }
        else {
            if (sp == 0) return;
            sp--;
            idx = st_idx[sp]; lo = st_lo[sp]; hi = st_hi[sp];
        }
    }
}

__device__ __forceinline__ uint4 ld_cg_u4(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

// one warp per list entry: the subtree's records and its leaf boxes [cl, ch] from the owner's buffers into
// mine.  Karras' numbering gives the ch - cl internal nodes of a subtree over [cl, ch] the indices cl..ch
// without ONE end: a left child is numbered by its last leaf (the root is node ch, index cl belongs to a
// node outside the subtree -- one that spans windows and is built locally), a right child by its first.
__global__ void __launch_bounds__(256)
halo_fetch_kernel(const uint2* __restrict__ list, const uint32_t* __restrict__ list_hdr, uint32_t list_cap,
                  const unsigned long long* __restrict__ wide_table, const unsigned long long* __restrict__ bounds_table,
                  const unsigned long long* __restrict__ ids_table, const uint32_t* __restrict__ win_edges,
                  uint32_t world, unsigned char* __restrict__ wide, unsigned char* __restrict__ bounds,
                  uint32_t* __restrict__ ids, uint32_t box_bytes, uint32_t n) {
    if (list_hdr[1]) {
        // more remote subtrees than the list holds: some would stay unfetched, and a walk into records that were
        // never written need not end.  Same remedy as in top_packed_kernel: a root that ends every walk of THIS
        // rank at once (the root is not among the records other ranks copy); the host raises when it next syncs.
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            uint4* rp = reinterpret_cast<uint4*>(wide);
            rp[0] = make_uint4(0x80000000u, 0x80000000u, 0xFFFFFFFFu, 0xFFFFFFFFu);
            rp[1] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
        }
        return;
    }
    const uint32_t count = min(list_hdr[0], list_cap);
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t leaf_start = n - 1;
    for (uint32_t e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); e < count; e += gridDim.x * (blockDim.x >> 5)) {
        uint2 ent = list[e];
        const bool left = (ent.y >> 31) != 0;
        ent.y &= 0x7fffffffu;
        const uint32_t w = ent.x >> 8;
        uint32_t owner = 0;
        while (owner + 1 < world && w >= win_edges[owner + 1]) owner++;
        const unsigned char* src_w = reinterpret_cast<const unsigned char*>(wide_table[owner]);
        const unsigned char* src_b = reinterpret_cast<const unsigned char*>(bounds_table[owner]);
        // records: 32 bytes each = 2 x 16; leaf boxes: box_bytes each
        const uint32_t nrec16 = (ent.y - ent.x) * 2, rec0 = ent.x + (left ? 1u : 0u);
        for (uint32_t k = lane; k < nrec16; k += 32) {
            const size_t off = (size_t)rec0 * QNODE_BYTES + (size_t)k * 16;
            *reinterpret_cast<uint4*>(wide + off) = ld_cg_u4(reinterpret_cast<const uint4*>(src_w + off));
        }
        const uint32_t nbox16 = (ent.y - ent.x + 1) * (box_bytes / 16);
        for (uint32_t k = lane; k < nbox16; k += 32) {
            const size_t off = (size_t)(leaf_start + ent.x) * box_bytes + (size_t)k * 16;
            *reinterpret_cast<uint4*>(bounds + off) = ld_cg_u4(reinterpret_cast<const uint4*>(src_b + off));
        }
        if (ids_table) {   // sharded sort: the sphere ids of those leaves are the owner's to give, too
            const uint32_t* src_i = reinterpret_cast<const uint32_t*>(ids_table[owner]);
            for (uint32_t k = ent.x + lane; k <= ent.y; k += 32) ids[k] = ld_relaxed_u32(src_i + k);
        }
    }
}

// ---- the nodes above the windows, two levels (sharded build).  LOCAL: a rank climbs from the roots its own
// windows handed over, as long as the parent still lies inside its own range of leaves [leaf_begin, leaf_end) --
// known from the node alone, like the window test: the parent reaches past the range's end iff the code just
// beyond it shares at least the parent's prefix.  Where it does, the subtree is one of the rank's (few) RANGE
// ROOTS and goes into the exchange; every node finished on the way is listed there by index.  GLOBAL
// (top_packed_kernel over the gathered range roots): the handful of nodes that span ranks, on every rank.
template <typename T>
__global__ void __launch_bounds__(256)
top_local_kernel(const uint32_t* __restrict__ codes, col_node* __restrict__ nodes, unsigned char* __restrict__ wide,
                 uint32_t* __restrict__ flags, T* __restrict__ bounds,
                 const T* __restrict__ scene, uint32_t n, uint32_t leaf_begin, uint32_t leaf_end,
                 const unsigned char* __restrict__ probe, uint32_t probe_cap, unsigned char* __restrict__ xchg,
                 uint32_t span_cap) {
    // one thread per root of the probe list (dense: one per hand-over SLOT left 98 % of the lanes idle and the
    // few busy ones spread over 25 x as many warps -- 2.07 ms instead of 0.5 at 2^26 own leaves)
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t* hdr = reinterpret_cast<uint32_t*>(xchg);
    const uint32_t* probe_hdr = reinterpret_cast<const uint32_t*>(probe);
    // a probe list that overflowed leaves the halo incomplete: everybody has to know (see shard_exchange_overflowed)
    if (j == 0 && probe_hdr[1]) hdr[1] = 1u;
    if (j >= min(probe_hdr[0], probe_cap)) return;
    const PackedItem<T> root = reinterpret_cast<const PackedItem<T>*>(probe + PACK_HEADER)[j];
    uint32_t self = root.self;
    const Quant<T> qz = make_quant(scene);
    const uint32_t leaf_start = n - 1;
    uint32_t lo = root.lo, hi = root.hi;
    Box<T> box;
#pragma unroll
    for (int c = 0; c < 3; c++) { box.lo[c] = root.box[c]; box.hi[c] = root.box[3 + c]; }
    uint3 grid = quantise_box(qz, box.lo, box.hi);
    uint32_t code_lo = root.code[1], code_hi = root.code[2];
    int dl = lo > 0 ? key_delta(lo - 1, root.code[0], lo, code_lo) : -1;
    int dr = hi < leaf_start ? key_delta(hi, code_hi, hi + 1, root.code[3]) : -1;
    const uint32_t code_before = leaf_begin > 0 ? __ldg(codes + leaf_begin - 1) : 0u;
    const uint32_t code_after = leaf_end < n ? __ldg(codes + leaf_end) : 0u;
    uint32_t* span = reinterpret_cast<uint32_t*>(xchg + xchg_span_offset<T>());
    while (true) {
        const bool is_left = dr > dl;
        const bool inside = is_left ? (leaf_end >= n || key_delta(hi, code_hi, leaf_end, code_after) < dr)
                                    : (leaf_begin == 0 || key_delta(leaf_begin - 1, code_before, lo, code_lo) < dl);
        if (!inside) {   // a range root
            const uint32_t k = atomicAdd(hdr, 1u);
            if (k >= XCHG_ROOTS) { hdr[1] = 1u; return; }
            PackedItem<T> it;
            it.self = self; it.lo = lo; it.hi = hi; it.pad = 0;
            it.code[0] = lo > 0 ? __ldg(codes + lo - 1) : 0u;
            it.code[1] = code_lo;
            it.code[2] = code_hi;
            it.code[3] = hi < leaf_start ? __ldg(codes + hi + 1) : 0u;
#pragma unroll
            for (int c = 0; c < 3; c++) { it.box[c] = box.lo[c]; it.box[3 + c] = box.hi[c]; }
            reinterpret_cast<PackedItem<T>*>(xchg + XCHG_HEADER)[k] = it;
            return;
        }
        const uint32_t split = is_left ? hi : lo - 1;
        const uint32_t far_end = exchange_acq_rel(flags + split, is_left ? lo : hi);
        if (far_end == NO_ARRIVAL) return;  // first to arrive: the sibling carries on
        uint32_t sibling;
        if (is_left) {
            sibling = far_end == split + 1 ? leaf_start + split + 1 : split + 1;
            hi = far_end;
            code_hi = __ldg(codes + hi);
            dr = hi < leaf_start ? key_delta(hi, code_hi, hi + 1, __ldg(codes + hi + 1)) : -1;
        } else {
            sibling = far_end == split ? leaf_start + split : split;
            lo = far_end;
            code_lo = __ldg(codes + lo);
            dl = lo > 0 ? key_delta(lo - 1, __ldg(codes + lo - 1), lo, code_lo) : -1;
        }
        const Box<T> other = load_box_cg(bounds, sibling);
        const uint3 other_grid = quantise_box(qz, other.lo, other.hi);
        self = finish_node(nodes, wide, bounds, leaf_start, lo, hi, dl, dr, is_left, self, sibling, box, other, grid, other_grid);
        {   // one slot of the list per finished node; the lanes that got here together take theirs with ONE atomic
            // (a counter bumped once per node by the whole grid was a third of this kernel's time)
            const unsigned m = __activemask();
            const unsigned lane = lane_id(), leader = (unsigned)__ffs(m) - 1u;
            uint32_t base = 0;
            if (lane == leader) base = atomicAdd(hdr + 2, (uint32_t)__popc(m));
            base = __shfl_sync(m, base, leader);
            const uint32_t k = base + (uint32_t)__popc(m & lanemask_lt());
            if (k < span_cap) span[k] = self;
            else hdr[3] = 1u;
        }
        if (lo == 0 && hi == leaf_start) return;
    }
}

// the other ranks' spanning nodes (by their lists) -> my copy of the records, from the owners' HBM
template <typename T>
__global__ void __launch_bounds__(256)
fetch_spanning_kernel(const unsigned char* __restrict__ gathered, size_t stride, uint32_t span_cap, uint32_t world,
                      uint32_t my_rank, const unsigned long long* __restrict__ wide_table, unsigned char* __restrict__ wide) {
    if (shard_exchange_overflowed(gathered, stride, world)) return;
    for (uint32_t r = 0; r < world; r++) {
        if (r == my_rank) continue;
        const unsigned char* theirs = gathered + r * stride;
        const uint32_t count = min(reinterpret_cast<const uint32_t*>(theirs)[2], span_cap);
        const uint32_t* list = reinterpret_cast<const uint32_t*>(theirs + xchg_span_offset<T>());
        const unsigned char* src = reinterpret_cast<const unsigned char*>(wide_table[r]);
        // two threads per 32-byte record
        for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < 2 * count; k += gridDim.x * blockDim.x) {
            const size_t off = (size_t)list[k >> 1] * QNODE_BYTES + (k & 1u) * 16;
            *reinterpret_cast<uint4*>(wide + off) = ld_cg_u4(reinterpret_cast<const uint4*>(src + off));
        }
    }
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

COL_API int col_build_refit_wide(const uint32_t* codes_sorted, const uint32_t* ids_sorted, void* nodes, void* wide,
                                 uint32_t* flags, void* bounds, const void* coords4, const void* radii,
                                 const void* scene2x4, int dtype, uint32_t n, void* stream) {
    if (!codes_sorted || !ids_sorted || !nodes || !wide || !flags || !bounds || !coords4 || !radii || !scene2x4)
        return COL_ERR_BAD_ARG;
    if (scene2x4 == bounds) return COL_ERR_BAD_ARG;  // slot 0 of the Bound array is overwritten with the root box
    if (n == 0) return 0;
    if (n > 0x80000000u) return COL_ERR_TOO_LARGE;
    cudaStream_t s = col_stream(stream);
    // flags: [0, n-1) arrival words of the cross-window splits, [n, 2n) hand-over slots of pass 1.
    // Two passes pay off once the scene is many waves of blocks deep (measured: 2^20 spheres
    // 0.146 vs 0.151 ms in one launch, 2^23 spheres 0.92 vs 0.83 ms in two).
    const bool two_pass = n >= COL_BR_TWO_PASS_MIN_N;
    COL_CUDA(cudaMemsetAsync(flags, 0xFF, sizeof(uint32_t) * (two_pass ? 2 * (size_t)n : (size_t)n), s));
    const unsigned blocks = col_blocks(n, BR_WINDOW);
    uint32_t* items = flags + n;
    // (Round 2, measured and dropped: the window pass done in barrier-separated ROUNDS with the merges of a
    //  round re-packed into full warps -- 36.8 M instead of 57 M warp instructions at 2^20 leaves, but 10.6
    //  rounds per window and 40 % of the stall samples behind the round barrier: 0.173 vs 0.148 ms, 2.01 vs
    //  1.62 ms at configs[3]; profiles/r2h_ncu_build_rounds_summary.txt.)
#define COL_BUILD_REFIT(T)                                                                                          \
    do {                                                                                                            \
        if (two_pass) {                                                                                             \
            build_refit_kernel<T, true><<<blocks, BR_WINDOW, 0, s>>>(                                       \
                codes_sorted, ids_sorted, (col_node*)nodes, (unsigned char*)wide, flags, items, (T*)bounds,         \
                (const T*)coords4, (const T*)radii, (const T*)scene2x4, n, 0u);                                     \
            build_refit_top_kernel<T><<<col_blocks(n, 256), 256, 0, s>>>(                                           \
                codes_sorted, (col_node*)nodes, (unsigned char*)wide, flags, items, (T*)bounds, (const T*)scene2x4, n); \
        } else {                                                                                                    \
            build_refit_kernel<T, false><<<blocks, BR_WINDOW, 0, s>>>(                                      \
                codes_sorted, ids_sorted, (col_node*)nodes, (unsigned char*)wide, flags, items, (T*)bounds,         \
                (const T*)coords4, (const T*)radii, (const T*)scene2x4, n, 0u);                                     \
        }                                                                                                           \
    } while (0)
    if (dtype == COL_F32) COL_BUILD_REFIT(float);
    else if (dtype == COL_F64) COL_BUILD_REFIT(double);
    else return COL_ERR_UNSUPPORTED;
#undef COL_BUILD_REFIT
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

// ---------------------------------------------------------------- sharded build: entry points
COL_API size_t col_shard_packed_bytes(uint32_t cap, int dtype) {
    return PACK_HEADER + (size_t)cap * (dtype == COL_F64 ? sizeof(PackedItem<double>) : sizeof(PackedItem<float>));
}

COL_API size_t col_shard_list_bytes(uint32_t list_cap) { return 16 + (size_t)list_cap * sizeof(uint2); }

COL_API int col_shard_build_windows(const uint32_t* codes_sorted, const uint32_t* ids_sorted, void* nodes, void* wide,
                                    uint32_t* flags, void* bounds, const void* coords4, const void* radii,
                                    const void* scene2x4, int dtype, uint32_t n, uint32_t win_begin, uint32_t win_end,
                                    void* packed, uint32_t cap, void* stream) {
    if (!codes_sorted || !ids_sorted || !nodes || !wide || !flags || !bounds || !coords4 || !radii || !scene2x4 || !packed)
        return COL_ERR_BAD_ARG;
    if (dtype != COL_F32 && dtype != COL_F64) return COL_ERR_UNSUPPORTED;
    const uint32_t windows = col_blocks(n, BR_WINDOW);
    if (n < 2 || win_begin > win_end || win_end > windows || cap == 0) return COL_ERR_BAD_ARG;
    cudaStream_t s = col_stream(stream);
    // flags: [0, n-1) arrival words of the cross-window splits, [n, 2n) hand-over slots / fetch marks
    COL_CUDA(cudaMemsetAsync(flags, 0xFF, sizeof(uint32_t) * 2 * (size_t)n, s));
    COL_CUDA(cudaMemsetAsync(packed, 0, PACK_HEADER, s));
    if (win_begin == win_end) return 0;
    uint32_t* items = flags + n;
    const uint32_t leaf_begin = win_begin * BR_WINDOW, leaf_end = min(win_end * BR_WINDOW, n);
    const unsigned blocks = win_end - win_begin, pblocks = col_blocks(leaf_end - leaf_begin, 256);
    if (dtype == COL_F32) {
        build_refit_kernel<float, true><<<blocks, BR_WINDOW, 0, s>>>(
            codes_sorted, ids_sorted, (col_node*)nodes, (unsigned char*)wide, flags, items, (float*)bounds,
            (const float*)coords4, (const float*)radii, (const float*)scene2x4, n, win_begin);
        pack_items_kernel<float><<<pblocks, 256, 0, s>>>(items, (const col_node*)nodes, (const float*)bounds, codes_sorted,
                                                         (unsigned char*)packed, cap, n, leaf_begin, leaf_end);
    } else {
        build_refit_kernel<double, true><<<blocks, BR_WINDOW, 0, s>>>(
            codes_sorted, ids_sorted, (col_node*)nodes, (unsigned char*)wide, flags, items, (double*)bounds,
            (const double*)coords4, (const double*)radii, (const double*)scene2x4, n, win_begin);
        pack_items_kernel<double><<<pblocks, 256, 0, s>>>(items, (const col_node*)nodes, (const double*)bounds, codes_sorted,
                                                          (unsigned char*)packed, cap, n, leaf_begin, leaf_end);
    }
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API size_t col_shard_exchange_bytes(uint32_t span_cap, int dtype) {
    const size_t off = dtype == COL_F64 ? xchg_span_offset<double>() : xchg_span_offset<float>();
    return (off + (size_t)span_cap * sizeof(uint32_t) + 15) & ~(size_t)15;
}

COL_API int col_shard_top_local(const uint32_t* codes_sorted, void* nodes, void* wide, uint32_t* flags, void* bounds,
                                const void* scene2x4, int dtype, uint32_t n, uint32_t leaf_begin, uint32_t leaf_end,
                                const void* probe, uint32_t probe_cap, void* exchange, uint32_t span_cap, void* stream) {
    if (!codes_sorted || !nodes || !wide || !flags || !bounds || !scene2x4 || !probe || !exchange || span_cap == 0 ||
        probe_cap == 0 || leaf_begin > leaf_end || leaf_end > n)
        return COL_ERR_BAD_ARG;
    if (dtype != COL_F32 && dtype != COL_F64) return COL_ERR_UNSUPPORTED;
    cudaStream_t s = col_stream(stream);
    COL_CUDA(cudaMemsetAsync(exchange, 0, XCHG_HEADER, s));
    if (leaf_begin == leaf_end) return 0;
    const unsigned blocks = col_blocks(probe_cap, 256);
    if (dtype == COL_F32)
        top_local_kernel<float><<<blocks, 256, 0, s>>>(codes_sorted, (col_node*)nodes, (unsigned char*)wide, flags,
                                                       (float*)bounds, (const float*)scene2x4, n, leaf_begin, leaf_end,
                                                       (const unsigned char*)probe, probe_cap, (unsigned char*)exchange,
                                                       span_cap);
    else
        top_local_kernel<double><<<blocks, 256, 0, s>>>(codes_sorted, (col_node*)nodes, (unsigned char*)wide, flags,
                                                        (double*)bounds, (const double*)scene2x4, n, leaf_begin, leaf_end,
                                                        (const unsigned char*)probe, probe_cap, (unsigned char*)exchange,
                                                        span_cap);
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_shard_top(const void* gathered, uint32_t span_cap, uint32_t world, uint32_t my_rank,
                          uint32_t* codes_sorted, void* nodes, void* wide, uint32_t* flags, void* bounds,
                          const void* scene2x4, int dtype, uint32_t n, const uint64_t* wide_table, void* stream) {
    if (!gathered || !codes_sorted || !nodes || !wide || !flags || !bounds || !scene2x4 || !wide_table || span_cap == 0 ||
        world == 0 || my_rank >= world)
        return COL_ERR_BAD_ARG;
    if (dtype != COL_F32 && dtype != COL_F64) return COL_ERR_UNSUPPORTED;
    cudaStream_t s = col_stream(stream);
    const size_t stride = col_shard_exchange_bytes(span_cap, dtype);
    const unsigned blocks = col_blocks((uint64_t)XCHG_ROOTS * world, 256);
    // (the same values the replicated sort left there; the only ones this rank has after the sharded sort)
    if (dtype == COL_F32) {
        scatter_root_codes_kernel<float><<<blocks, 256, 0, s>>>((const unsigned char*)gathered, stride, XCHG_ROOTS, world,
                                                                my_rank, codes_sorted, n);
        top_packed_kernel<float><<<blocks, 256, 0, s>>>((const unsigned char*)gathered, stride, XCHG_ROOTS, world, my_rank,
                                                        codes_sorted, (col_node*)nodes, (unsigned char*)wide, flags,
                                                        (float*)bounds, (const float*)scene2x4, n);
        fetch_spanning_kernel<float><<<COL_NUM_SMS * 8, 256, 0, s>>>((const unsigned char*)gathered, stride, span_cap, world,
                                                                     my_rank, (const unsigned long long*)wide_table,
                                                                     (unsigned char*)wide);
    } else {
        scatter_root_codes_kernel<double><<<blocks, 256, 0, s>>>((const unsigned char*)gathered, stride, XCHG_ROOTS, world,
                                                                 my_rank, codes_sorted, n);
        top_packed_kernel<double><<<blocks, 256, 0, s>>>((const unsigned char*)gathered, stride, XCHG_ROOTS, world, my_rank,
                                                         codes_sorted, (col_node*)nodes, (unsigned char*)wide, flags,
                                                         (double*)bounds, (const double*)scene2x4, n);
        fetch_spanning_kernel<double><<<COL_NUM_SMS * 8, 256, 0, s>>>((const unsigned char*)gathered, stride, span_cap, world,
                                                                      my_rank, (const unsigned long long*)wide_table,
                                                                      (unsigned char*)wide);
    }
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_shard_halo(const void* probe, uint32_t cap, const void* gathered, uint32_t span_cap, uint32_t world,
                           uint32_t my_rank, void* wide, void* bounds,
                           const void* scene2x4, int dtype, uint32_t* flags, void* list, uint32_t list_cap,
                           const uint64_t* wide_table, const uint64_t* bounds_table, const uint32_t* win_edges,
                           uint32_t n, uint32_t leaf_begin, uint32_t leaf_end, const uint64_t* ids_table,
                           uint32_t* ids_sorted, void* stream) {
    if (!probe || !gathered || !wide || !bounds || !scene2x4 || !flags || !list || !wide_table || !bounds_table || !win_edges ||
        cap == 0 || span_cap == 0 || list_cap == 0 || my_rank >= world || (ids_table && !ids_sorted))
        return COL_ERR_BAD_ARG;
    if (dtype != COL_F32 && dtype != COL_F64) return COL_ERR_UNSUPPORTED;
    cudaStream_t s = col_stream(stream);
    COL_CUDA(cudaMemsetAsync(list, 0, 16, s));
    uint32_t* hdr = reinterpret_cast<uint32_t*>(list);
    uint2* entries = reinterpret_cast<uint2*>(reinterpret_cast<unsigned char*>(list) + 16);
    const size_t stride = col_shard_exchange_bytes(span_cap, dtype);
    const unsigned char* mine = reinterpret_cast<const unsigned char*>(probe);
    uint32_t* asked = flags + n;
    const unsigned pblocks = col_blocks(cap, 256);
    if (dtype == COL_F32)
        halo_probe_kernel<float><<<pblocks, 256, 0, s>>>((const unsigned char*)gathered, stride, world, mine, cap,
                                                         (const unsigned char*)wide, (const float*)scene2x4, asked,
                                                         entries, hdr, list_cap, n, leaf_begin, leaf_end);
    else
        halo_probe_kernel<double><<<pblocks, 256, 0, s>>>((const unsigned char*)gathered, stride, world, mine, cap,
                                                          (const unsigned char*)wide, (const double*)scene2x4,
                                                          asked, entries, hdr, list_cap, n, leaf_begin, leaf_end);
    COL_CHECK_LAUNCH();
    halo_fetch_kernel<<<COL_NUM_SMS * 8, 256, 0, s>>>(entries, hdr, list_cap, (const unsigned long long*)wide_table,
                                                      (const unsigned long long*)bounds_table,
                                                      (const unsigned long long*)ids_table, win_edges, world,
                                                      (unsigned char*)wide, (unsigned char*)bounds, ids_sorted,
                                                      dtype == COL_F64 ? 64u : 32u, n);
    COL_CHECK_LAUNCH();
    return 0;
}
