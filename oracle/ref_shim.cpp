// ref_shim.cpp -- TEST INFRASTRUCTURE.  Compiles the reference's OWN barrier-free
// OpenCL C kernels (collision.cl, index.cl, offset.cl) as C++ by #including the
// .cl text where it lies under /root/reference (never copied into this repo)
// behind a small OpenCL-C compatibility layer, and exports one C entry point per
// kernel that runs the work-items of a 1-D NDRange sequentially in gid order.
//
// Used to validate oracle/collision_oracle.c against the real reference source
// (tests/test_oracle_vs_ref.py) and to generate tests/golden/ fixtures.  The
// barrier-using kernels (radix.cl, scan.cl, reduce.cl) cannot be run by a
// sequential work-item loop and are not included; their results are fully
// specified by stable-sort / cumsum / min-max and pinned by the reference's
// golden vectors instead.
//
// Build: see oracle/Makefile (g++ -DDTYPE_IS_FLOAT|-DDTYPE_IS_DOUBLE -I<ref>/collision).
#include <climits>
#include <cmath>
#include <cstddef>
#include <cstdint>

#if defined(DTYPE_IS_DOUBLE)
#define DTYPE double
#else
#define DTYPE float
#endif

namespace clshim {

static thread_local size_t g_gid = 0;
static inline size_t get_global_id(int) { return g_gid; }

template <typename T>
struct alignas(4 * sizeof(T)) vec3 {
    T x, y, z, pad;
};
struct int3 {
    int x, y, z;
};
typedef vec3<float> float3;
typedef vec3<double> double3;

template <typename T> static inline vec3<T> operator-(vec3<T> a, vec3<T> b) { return {a.x - b.x, a.y - b.y, a.z - b.z, 0}; }
template <typename T> static inline vec3<T> operator/(vec3<T> a, vec3<T> b) { return {a.x / b.x, a.y / b.y, a.z / b.z, 0}; }
template <typename T> static inline vec3<T> operator*(vec3<T> a, T s) { return {a.x * s, a.y * s, a.z * s, 0}; }
template <typename T> static inline vec3<T> operator-(vec3<T> a, T s) { return {a.x - s, a.y - s, a.z - s, 0}; }
template <typename T> static inline vec3<T> operator+(vec3<T> a, T s) { return {a.x + s, a.y + s, a.z + s, 0}; }
// OpenCL vector relational operators give -1 (all bits) for true.
template <typename T> static inline int3 operator>(vec3<T> a, vec3<T> b) { return {-(a.x > b.x), -(a.y > b.y), -(a.z > b.z)}; }
template <typename T> static inline int3 operator<(vec3<T> a, vec3<T> b) { return {-(a.x < b.x), -(a.y < b.y), -(a.z < b.z)}; }
static inline int3 operator&(int3 a, int3 b) { return {a.x & b.x, a.y & b.y, a.z & b.z}; }
static inline bool all(int3 v) { return (v.x < 0) && (v.y < 0) && (v.z < 0); }

static inline float sfmin(float a, float b) { return fminf(a, b); }
static inline float sfmax(float a, float b) { return fmaxf(a, b); }
static inline double sfmin(double a, double b) { return fmin(a, b); }
static inline double sfmax(double a, double b) { return fmax(a, b); }

// OpenCL: clamp(x, lo, hi) == fmin(fmax(x, lo), hi)
template <typename T, typename L, typename H>
static inline vec3<T> clamp(vec3<T> v, L lo, H hi) {
    return {sfmin(sfmax(v.x, (T)lo), (T)hi), sfmin(sfmax(v.y, (T)lo), (T)hi),
            sfmin(sfmax(v.z, (T)lo), (T)hi), 0};
}
template <typename T> static inline vec3<T> min(vec3<T> a, vec3<T> b) { return {sfmin(a.x, b.x), sfmin(a.y, b.y), sfmin(a.z, b.z), 0}; }
template <typename T> static inline vec3<T> max(vec3<T> a, vec3<T> b) { return {sfmax(a.x, b.x), sfmax(a.y, b.y), sfmax(a.z, b.z), 0}; }
static inline unsigned int min(unsigned int a, unsigned int b) { return a < b ? a : b; }
static inline unsigned int max(unsigned int a, unsigned int b) { return a > b ? a : b; }
static inline int clz(unsigned int v) { return v ? __builtin_clz(v) : 32; }

static inline unsigned int atomic_inc(unsigned int *p) { return __atomic_fetch_add(p, 1u, __ATOMIC_SEQ_CST); }
#define CLK_GLOBAL_MEM_FENCE 1
static inline void mem_fence(int) { __atomic_thread_fence(__ATOMIC_SEQ_CST); }

#define kernel
#define global
#define local

namespace col {
#include "collision.cl"
}  // namespace col

namespace idx_u32_u32 {
#define VALUE_TYPE uint32_t
#define INDEX_TYPE uint32_t
#include "index.cl"
#undef VALUE_TYPE
#undef INDEX_TYPE
}  // namespace idx_u32_u32

namespace idx_f64_u64 {
#define VALUE_TYPE double
#define INDEX_TYPE uint64_t
#include "index.cl"
#undef VALUE_TYPE
#undef INDEX_TYPE
}  // namespace idx_f64_u64

namespace off_u32_u32 {
#define VALUE_TYPE uint32_t
#define OFFSET_TYPE uint32_t
#include "offset.cl"
#undef VALUE_TYPE
#undef OFFSET_TYPE
}  // namespace off_u32_u32

namespace off_u64_u64 {
#define VALUE_TYPE uint64_t
#define OFFSET_TYPE uint64_t
#include "offset.cl"
#undef VALUE_TYPE
#undef OFFSET_TYPE
}  // namespace off_u64_u64

#undef kernel
#undef global
#undef local

}  // namespace clshim

#define REF_API extern "C" __attribute__((visibility("default")))
#define FOR_GID(gsz) for (clshim::g_gid = 0; clshim::g_gid < (size_t)(gsz); clshim::g_gid++)

using clshim::col::Bound;
using clshim::col::Node;
typedef clshim::vec3<DTYPE> V3;

REF_API int ref_dtype_bytes(void) { return (int)sizeof(DTYPE); }

REF_API void ref_range(size_t gsz, unsigned int *idxs) { FOR_GID(gsz) clshim::col::range(idxs); }

REF_API void ref_calculateCodes(size_t gsz, unsigned int *codes, const void *coords,
                                const void *range, unsigned int n) {
    FOR_GID(gsz) clshim::col::calculateCodes(codes, (const V3 *)coords, (const V3 *)range, n);
}

REF_API void ref_fillInternal(size_t gsz, void *nodes, const unsigned int *ids, unsigned int n) {
    FOR_GID(gsz) clshim::col::fillInternal((Node *)nodes, ids, n);
}

REF_API void ref_generateBVH(size_t gsz, const unsigned int *codes, void *nodes, unsigned int n) {
    FOR_GID(gsz) clshim::col::generateBVH(codes, (Node *)nodes, n);
}

REF_API void ref_leafBounds(size_t gsz, void *bounds, const void *coords, const void *radii,
                            const void *nodes, unsigned int n) {
    FOR_GID(gsz)
    clshim::col::leafBounds((Bound *)bounds, (const V3 *)coords, (const DTYPE *)radii,
                            (const Node *)nodes, n);
}

REF_API void ref_internalBounds(size_t gsz, void *bounds, unsigned int *flags, const void *nodes,
                                unsigned int n) {
    FOR_GID(gsz) clshim::col::internalBounds((Bound *)bounds, flags, (const Node *)nodes, n);
}

REF_API void ref_traverse(size_t gsz, unsigned int *collisions, unsigned int *next,
                          unsigned int n_collisions, const void *nodes, const void *bounds,
                          unsigned int n) {
    FOR_GID(gsz)
    clshim::col::traverse(collisions, next, n_collisions, (const Node *)nodes,
                          (const Bound *)bounds, n);
}

REF_API void ref_gather_u32_u32(size_t gsz, const uint32_t *in, const uint32_t *idx, uint32_t *out) {
    FOR_GID(gsz) clshim::idx_u32_u32::gather(in, idx, out);
}
REF_API void ref_scatter_u32_u32(size_t gsz, const uint32_t *in, const uint32_t *idx, uint32_t *out) {
    FOR_GID(gsz) clshim::idx_u32_u32::scatter(in, idx, out);
}
REF_API void ref_gather_f64_u64(size_t gsz, const double *in, const uint64_t *idx, double *out) {
    FOR_GID(gsz) clshim::idx_f64_u64::gather(in, idx, out);
}
REF_API void ref_scatter_f64_u64(size_t gsz, const double *in, const uint64_t *idx, double *out) {
    FOR_GID(gsz) clshim::idx_f64_u64::scatter(in, idx, out);
}
REF_API void ref_find_offsets_u32_u32(size_t gsz, const uint32_t *values, uint32_t *offsets) {
    FOR_GID(gsz) clshim::off_u32_u32::find_offsets(values, offsets);
}
REF_API void ref_find_offsets_u64_u64(size_t gsz, const uint64_t *values, uint64_t *offsets) {
    FOR_GID(gsz) clshim::off_u64_u64::find_offsets(values, offsets);
}
