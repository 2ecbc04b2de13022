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
                   uint32_t n) {
    __shared__ T s_box[256][2][6];          // [split & 255][which child][lo xyz, hi xyz]
    __shared__ uint32_t s_grid[256][2][3];  // ... and its grid box
    __shared__ uint32_t s_far[256];         // arrival word of the in-window splits
    __shared__ uint32_t s_code[258];        // codes of positions first-1 .. last+1
    __shared__ uint32_t s_items;            // subtree roots handed on so far (HAND_OVER)
    const uint32_t first = blockIdx.x * blockDim.x;
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
                (const T*)coords4, (const T*)radii, (const T*)scene2x4, n);                                         \
            build_refit_top_kernel<T><<<col_blocks(n, 256), 256, 0, s>>>(                                           \
                codes_sorted, (col_node*)nodes, (unsigned char*)wide, flags, items, (T*)bounds, (const T*)scene2x4, n); \
        } else {                                                                                                    \
            build_refit_kernel<T, false><<<blocks, BR_WINDOW, 0, s>>>(                                      \
                codes_sorted, ids_sorted, (col_node*)nodes, (unsigned char*)wide, flags, items, (T*)bounds,         \
                (const T*)coords4, (const T*)radii, (const T*)scene2x4, n);                                         \
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
