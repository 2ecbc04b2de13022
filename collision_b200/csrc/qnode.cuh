// qnode.cuh -- the 32-byte traversal record and its conservative quantisation.
//
// One record per internal node (written by tree.cu + refit.cu, read by traverse_wide.cu):
//   w0 = child a, w1 = child b: an internal node as its index, a leaf as 0x80000000 | its sorted
//        position (so "is it a leaf", "its right edge" and "what to report" need no n),
//   w2..w4 = box of child a, w5..w7 = box of child b, each as 6 x u16
//             {lo.x lo.y | lo.z ~hi.x | ~hi.y ~hi.z}          (~v = 65535 - v)
// A visit is ONE 256-bit load.  The query carries {hi.x hi.y | hi.z ~lo.x | ~lo.y ~lo.z}, so the
// six comparisons of the box test all read "record half-word <= query half-word" and are
// evaluated two at a time: max.u16x2(record, query) == query  (VIMNMX.U16x2 + ISETP).  The u16 coordinates live on a 65536^3 grid spanning the scene
// bounds (of the centres; anything outside clamps to the border cells) and are rounded
// OUTWARD (floor for lo, ceil for hi), the query box likewise, and compared with >= / <=:
// an exact overlap always implies a grid overlap, so an internal node is never pruned
// wrongly; at worst a few extra nodes are visited.  The decision that matters -- is a LEAF
// reported -- is always taken on the exact float boxes, exactly as collision.cl:164-166
// does; since every ancestor's exact box contains the leaf's, the reported set and its DFS
// order are identical to the reference walk's.
//
// Right edges need no storage: child a of node X is a left child, so in Karras numbering
// its index IS its right edge (internal), as is its position (leaf): w0 & 0x7fffffff either
// way; child b's right edge is X's own, which was already checked (> q) when X was entered.
#pragma once
#include "common.cuh"

template <typename T> struct Quant {
    T lo[3], scale[3];
};

template <typename T>
__device__ __forceinline__ Quant<T> make_quant(const T* __restrict__ scene /* {min[4], max[4]} */) {
    Quant<T> qz;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const T lo = scene[k], ext = scene[4 + k] - scene[k];
        qz.lo[k] = lo;
        qz.scale[k] = ext > (T)0 ? (T)65535 / ext : (T)0;  // degenerate / NaN extent: everything maps to cell 0
    }
    return qz;
}

// monotone non-decreasing in x (subtract and multiply by a non-negative constant, each
// correctly rounded, then clamp); NaN -> 0
__device__ __forceinline__ float grid_coord(float x, float lo, float scale) {
    return fminf(fmaxf(__fmul_rn(__fsub_rn(x, lo), scale), 0.0f), 65535.0f);
}
__device__ __forceinline__ double grid_coord(double x, double lo, double scale) {
    return fmin(fmax(__dmul_rn(__dsub_rn(x, lo), scale), 0.0), 65535.0);
}
__device__ __forceinline__ uint32_t grid_floor(float t) { return (uint32_t)__float2int_rd(t); }
__device__ __forceinline__ uint32_t grid_ceil(float t) { return (uint32_t)__float2int_ru(t); }
__device__ __forceinline__ uint32_t grid_floor(double t) { return (uint32_t)__double2int_rd(t); }
__device__ __forceinline__ uint32_t grid_ceil(double t) { return (uint32_t)__double2int_ru(t); }

// box {lo[3], hi[3]} -> 3 words of 6 x u16 {lo.x lo.y | lo.z ~hi.x | ~hi.y ~hi.z}, rounded outward
template <typename T>
__device__ __forceinline__ uint3 quantise_box(const Quant<T>& qz, const T (&lo)[3], const T (&hi)[3]) {
    uint32_t v[6];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        v[k] = grid_floor(grid_coord(lo[k], qz.lo[k], qz.scale[k]));
        v[3 + k] = 65535u - grid_ceil(grid_coord(hi[k], qz.lo[k], qz.scale[k]));
    }
    return make_uint3(v[0] | (v[1] << 16), v[2] | (v[3] << 16), v[4] | (v[5] << 16));
}
// the query's side of the same test: {hi.x hi.y | hi.z ~lo.x | ~lo.y ~lo.z}, rounded outward
template <typename T>
__device__ __forceinline__ uint3 quantise_query(const Quant<T>& qz, const T (&lo)[3], const T (&hi)[3]) {
    uint32_t v[6];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        v[k] = grid_ceil(grid_coord(hi[k], qz.lo[k], qz.scale[k]));
        v[3 + k] = 65535u - grid_floor(grid_coord(lo[k], qz.lo[k], qz.scale[k]));
    }
    return make_uint3(v[0] | (v[1] << 16), v[2] | (v[3] << 16), v[4] | (v[5] << 16));
}
__device__ __forceinline__ uint32_t max_u16x2(uint32_t a, uint32_t b) {
    uint32_t d;
    asm("max.u16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t min_u16x2(uint32_t a, uint32_t b) {
    uint32_t d;
    asm("min.u16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
// grid box of the union of two boxes.  grid_coord, floor and ceil are monotone, so they commute
// with the min / max of the exact merge: the union's lo cells are the smaller lo cells and its
// ~hi cells the smaller ~hi cells -- one packed minimum per word, no float arithmetic.
__device__ __forceinline__ uint3 grid_union(const uint3& a, const uint3& b) {
    return make_uint3(min_u16x2(a.x, b.x), min_u16x2(a.y, b.y), min_u16x2(a.z, b.z));
}
// every half-word of the record's box <= the query's  <=>  the grid boxes overlap
__device__ __forceinline__ bool grid_overlap(const uint3& query, uint32_t r0, uint32_t r1, uint32_t r2) {
    return (max_u16x2(r0, query.x) == query.x) & (max_u16x2(r1, query.y) == query.y) &
           (max_u16x2(r2, query.z) == query.z);
}

constexpr uint32_t QNODE_BYTES = 32;
