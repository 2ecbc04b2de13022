/*
 * collision_oracle.c -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * A plain-C restatement of the arithmetic of kwohlfahrt/collision's LBVH
 * sphere-collision path, used ONLY as the checker in tests/, in
 * __graft_entry__.smoke() and as bench.py's cpu_baseline / --impl reference
 * leg.  The shipped package (collision_b200/) never links, imports or calls
 * anything in this directory.
 *
 * Parity status: PINNED.  Every function here is checked in tests/ against
 * the golden vectors of the reference's own test-suite
 * and, in the build container, against the reference's barrier-free OpenCL
 * kernels compiled as C++ from /root/reference (oracle/_ref, see
 * oracle/ref_shim.cpp).
 *
 * Each function cites the reference file:line it follows (paths relative to
 * the reference checkout).
 *
 * Build: gcc -O3 -march=native -fopenmp -ffp-contract=off -fno-fast-math -shared
 * -fPIC (oracle/Makefile; rebuilt per host CPU, see oracle/__init__.py:build).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

typedef struct {
    uint32_t parent;
    uint32_t right_edge;
    uint32_t data[2]; /* leaf: data[0] = id ; internal: children[2] */
} orc_node;           /* collision/collision.cl:42-53, collision/collision.py:9 */

ORC_API int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* Launchers such as torchrun export OMP_NUM_THREADS=1; the timed baseline sets the
 * thread count it reports explicitly. */
ORC_API void orc_set_num_threads(int nt) {
#ifdef _OPENMP
    if (nt > 0) omp_set_num_threads(nt);
#else
    (void)nt;
#endif
}

/* ------------------------------------------------------------------ bounds
 * collision/reduce.cl:5-58 with the accumulator list of collision/bounds.py:5
 * ([(INFINITY, min), (-INFINITY, max)]).  `lanes` is the slot width in scalars
 * (1 for scalar dtypes, 4 for 3- and 4-vectors: collision/misc.py:62-68).
 * out = {min[lanes], max[lanes]}.  Every lane is reduced independently.
 *
 * ngroups/group_size == 0: true min/max.  Otherwise the launch geometry of
 * collision/reduce.py:62-76 is emulated, including the bounds2 tree
 * (reduce.cl:42-52: o = ngroups/2; o > 0; o /= 2) that silently drops tail
 * partials when ngroups is not a power of two.
 */
#define DEF_BOUNDS(NAME, T, FMIN, FMAX)                                              \
    ORC_API void NAME(const T *vals, uint64_t n, uint32_t lanes, T *out,             \
                      uint32_t ngroups, uint32_t group_size) {                       \
        if (ngroups == 0 || group_size == 0) {                                       \
            for (uint32_t l = 0; l < lanes; l++) {                                   \
                T lo = (T)INFINITY, hi = (T)-INFINITY;                               \
                for (uint64_t i = 0; i < n; i++) {                                   \
                    lo = FMIN(lo, vals[i * lanes + l]);                              \
                    hi = FMAX(hi, vals[i * lanes + l]);                              \
                }                                                                    \
                out[l] = lo;                                                         \
                out[lanes + l] = hi;                                                 \
            }                                                                        \
            return;                                                                  \
        }                                                                            \
        uint64_t gsz = (uint64_t)ngroups * group_size;                               \
        T *part = (T *)malloc(sizeof(T) * 2 * lanes * ngroups);                      \
        T *scr = (T *)malloc(sizeof(T) * 2 * lanes * group_size);                    \
        for (uint32_t g = 0; g < ngroups; g++) {                                     \
            for (uint32_t t = 0; t < group_size; t++) {                              \
                for (uint32_t l = 0; l < lanes; l++) {                               \
                    T lo = (T)INFINITY, hi = (T)-INFINITY;                           \
                    for (uint64_t i = (uint64_t)g * group_size + t; i < n; i += gsz) { \
                        lo = FMIN(lo, vals[i * lanes + l]);                          \
                        hi = FMAX(hi, vals[i * lanes + l]);                          \
                    }                                                                \
                    scr[(t * 2 + 0) * lanes + l] = lo;                               \
                    scr[(t * 2 + 1) * lanes + l] = hi;                               \
                }                                                                    \
            }                                                                        \
            for (uint32_t o = group_size / 2; o > 0; o /= 2)                         \
                for (uint32_t t = 0; t < o; t++)                                     \
                    for (uint32_t l = 0; l < lanes; l++) {                           \
                        scr[(t * 2 + 0) * lanes + l] =                               \
                            FMIN(scr[(t * 2 + 0) * lanes + l],                       \
                                 scr[((t + o) * 2 + 0) * lanes + l]);                \
                        scr[(t * 2 + 1) * lanes + l] =                               \
                            FMAX(scr[(t * 2 + 1) * lanes + l],                       \
                                 scr[((t + o) * 2 + 1) * lanes + l]);                \
                    }                                                                \
            memcpy(part + (size_t)g * 2 * lanes, scr, sizeof(T) * 2 * lanes);        \
        }                                                                            \
        for (uint32_t o = ngroups / 2; o > 0; o /= 2)                                \
            for (uint32_t t = 0; t < o; t++)                                         \
                for (uint32_t l = 0; l < lanes; l++) {                               \
                    part[(t * 2 + 0) * lanes + l] =                                  \
                        FMIN(part[(t * 2 + 0) * lanes + l],                          \
                             part[((t + o) * 2 + 0) * lanes + l]);                   \
                    part[(t * 2 + 1) * lanes + l] =                                  \
                        FMAX(part[(t * 2 + 1) * lanes + l],                          \
                             part[((t + o) * 2 + 1) * lanes + l]);                   \
                }                                                                    \
        memcpy(out, part, sizeof(T) * 2 * lanes);                                    \
        free(part);                                                                  \
        free(scr);                                                                   \
    }

DEF_BOUNDS(orc_bounds_f32, float, fminf, fmaxf)
DEF_BOUNDS(orc_bounds_f64, double, fmin, fmax)

/* Multi-threaded true min/max over 4-lane slots (timed baseline only). */
ORC_API void orc_bounds_f32x4_mt(const float *vals, uint64_t n, float *out) {
    float lo[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
    float hi[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma omp parallel
    {
        float l[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
        float h[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma omp for nowait schedule(static)
        for (int64_t i = 0; i < (int64_t)n; i++)
            for (int k = 0; k < 4; k++) {
                l[k] = fminf(l[k], vals[i * 4 + k]);
                h[k] = fmaxf(h[k], vals[i * 4 + k]);
            }
#pragma omp critical
        for (int k = 0; k < 4; k++) {
            lo[k] = fminf(lo[k], l[k]);
            hi[k] = fmaxf(hi[k], h[k]);
        }
    }
    memcpy(out, lo, 16);
    memcpy(out + 4, hi, 16);
}

/* ------------------------------------------------------------------ Morton
 * collision/collision.cl:14-20 (expandBits) */
static inline uint32_t expand_bits(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

/* collision/collision.cl:22-40 (morton, calculateCodes).  coords are 4-lane
 * slots; range = {min[4], max[4]}.  IEEE sub / correctly rounded div / mul,
 * clamp == fmin(fmax(x, 0), scale) (OpenCL clamp), C truncating cast.
 * A NaN (0/0 on a degenerate axis) clamps to 0 through fmax -- outside the
 * reference's defined domain, fixed here so the CUDA path has something to
 * agree with. */
#define DEF_MORTON(NAME, T, FMIN, FMAX)                                      \
    ORC_API void NAME(uint32_t *codes, const T *coords, const T *range,      \
                      uint32_t n) {                                          \
        const T scale = (T)((1 << (sizeof(uint32_t) * 8 / 3)) - 1);          \
        _Pragma("omp parallel for schedule(static)")                         \
        for (int64_t i = 0; i < (int64_t)n; i++) {                           \
            uint32_t u[3];                                                   \
            for (int k = 0; k < 3; k++) {                                    \
                T p = (coords[i * 4 + k] - range[k]) /                       \
                      (range[4 + k] - range[k]);                             \
                p = p * scale;                                               \
                p = FMIN(FMAX(p, (T)0.0f), scale);                           \
                u[k] = (uint32_t)p;                                          \
            }                                                                \
            codes[i] = (expand_bits(u[0]) << 2) + (expand_bits(u[1]) << 1) + \
                       expand_bits(u[2]);                                    \
        }                                                                    \
    }

DEF_MORTON(orc_morton_f32, float, fminf, fmaxf)
DEF_MORTON(orc_morton_f64, double, fmin, fmax)

/* collision/collision.cl:8-10 (range) */
ORC_API void orc_range(uint32_t *ids, uint32_t n) {
    for (uint32_t i = 0; i < n; i++) ids[i] = i;
}

/* -------------------------------------------------------------------- sort
 * One LSD pass == stable partition by digit.  This is what block_sort +
 * prefix_sum + scatter compute together (collision/radix.cl:48-139,
 * collision/radix.py:140-157): the reference's tests define every pass as a
 * stable (mergesort) argsort by digit (tests/test_radix.py:150-182, :222-253).
 * key_bytes 4 or 8; value_bytes arbitrary (0 = keys only).
 */
static inline uint64_t load_key(const void *keys, uint64_t i, uint32_t key_bytes) {
    return key_bytes == 4 ? (uint64_t)((const uint32_t *)keys)[i]
                          : ((const uint64_t *)keys)[i];
}

ORC_API void orc_radix_pass(const void *keys_in, void *keys_out, const void *vals_in,
                            void *vals_out, uint64_t n, uint32_t key_bytes,
                            uint32_t value_bytes, uint32_t radix_bits, uint32_t pass) {
    const uint32_t H = 1u << radix_bits;
    const uint32_t shift = pass * radix_bits;
    uint64_t *count = (uint64_t *)calloc(H + 1, sizeof(uint64_t));
    for (uint64_t i = 0; i < n; i++)
        count[((load_key(keys_in, i, key_bytes) >> shift) & (H - 1)) + 1]++;
    for (uint32_t d = 0; d < H; d++) count[d + 1] += count[d];
    for (uint64_t i = 0; i < n; i++) {
        uint64_t k = load_key(keys_in, i, key_bytes);
        uint64_t dst = count[(k >> shift) & (H - 1)]++;
        memcpy((char *)keys_out + dst * key_bytes, (const char *)keys_in + i * key_bytes,
               key_bytes);
        if (value_bytes && vals_in && vals_out)
            memcpy((char *)vals_out + dst * value_bytes,
                   (const char *)vals_in + i * value_bytes, value_bytes);
    }
    free(count);
}

/* Full stable sort == all passes of collision/radix.py:140-170.  Implemented
 * as a multi-threaded LSD radix sort with 8-bit digits so the timed CPU
 * baseline is not handicapped; the result of ANY stable sort is identical
 * (tests/test_radix.py:252, :337-351; tests/test_radix_py.py:201).
 * keys_in/vals_in are not modified; scratch is allocated internally. */
ORC_API void orc_sort_pairs(const void *keys_in, void *keys_out, const void *vals_in,
                            void *vals_out, uint64_t n, uint32_t key_bytes,
                            uint32_t value_bytes) {
    const int have_vals = value_bytes && vals_in && vals_out;
    const uint32_t passes = key_bytes;
    char *kbuf[2], *vbuf[2] = {NULL, NULL};
    kbuf[0] = (char *)malloc((size_t)n * key_bytes + 8);
    kbuf[1] = (char *)malloc((size_t)n * key_bytes + 8);
    memcpy(kbuf[0], keys_in, (size_t)n * key_bytes);
    if (have_vals) {
        vbuf[0] = (char *)malloc((size_t)n * value_bytes + 8);
        vbuf[1] = (char *)malloc((size_t)n * value_bytes + 8);
        memcpy(vbuf[0], vals_in, (size_t)n * value_bytes);
    }
    int nt = orc_num_threads();
    uint64_t *hist = (uint64_t *)malloc(sizeof(uint64_t) * 256 * (size_t)nt);
    int cur = 0;
    for (uint32_t p = 0; p < passes; p++) {
        const uint32_t shift = p * 8;
        const char *ki = kbuf[cur], *vi = vbuf[cur];
        char *ko = kbuf[cur ^ 1], *vo = vbuf[cur ^ 1];
#pragma omp parallel num_threads(nt)
        {
#ifdef _OPENMP
            int t = omp_get_thread_num();
            int T = omp_get_num_threads();
#else
            int t = 0, T = 1;
#endif
            uint64_t lo = n * (uint64_t)t / T, hi = n * (uint64_t)(t + 1) / T;
            uint64_t *h = hist + 256 * (size_t)t;
            memset(h, 0, 256 * sizeof(uint64_t));
            for (uint64_t i = lo; i < hi; i++)
                h[(load_key(ki, i, key_bytes) >> shift) & 255]++;
#pragma omp barrier
#pragma omp single
            {
                uint64_t run = 0;
                for (int d = 0; d < 256; d++)
                    for (int tt = 0; tt < T; tt++) {
                        uint64_t c = hist[256 * (size_t)tt + d];
                        hist[256 * (size_t)tt + d] = run;
                        run += c;
                    }
            }
            for (uint64_t i = lo; i < hi; i++) {
                uint64_t k = load_key(ki, i, key_bytes);
                uint64_t dst = h[(k >> shift) & 255]++;
                if (key_bytes == 4) ((uint32_t *)ko)[dst] = (uint32_t)k;
                else ((uint64_t *)ko)[dst] = k;
                if (have_vals) {
                    if (value_bytes == 4) ((uint32_t *)vo)[dst] = ((const uint32_t *)vi)[i];
                    else memcpy(vo + dst * value_bytes, vi + i * value_bytes, value_bytes);
                }
            }
        }
        cur ^= 1;
    }
    memcpy(keys_out, kbuf[cur], (size_t)n * key_bytes);
    if (have_vals) memcpy(vals_out, vbuf[cur], (size_t)n * value_bytes);
    free(hist);
    free(kbuf[0]);
    free(kbuf[1]);
    free(vbuf[0]);
    free(vbuf[1]);
}

/* -------------------------------------------------------------------- scan
 * In-place exclusive prefix sum of uint32, wrapping mod 2^32
 * (collision/scan.py:75-112, collision/scan.cl:5-36,
 * collision/local_scan.cl:2-25; defined by tests/test_scan_py.py:57-64 as the
 * shifted cumsum). */
ORC_API void orc_exclusive_scan_u32(uint32_t *data, uint64_t n) {
    uint32_t run = 0;
    for (uint64_t i = 0; i < n; i++) {
        uint32_t v = data[i];
        data[i] = run;
        run += v;
    }
}

/* -------------------------------------------------------------------- tree
 * collision/collision.cl:55-63 (fillInternal -- fills the LEAVES). */
ORC_API void orc_fill_leaves(orc_node *nodes, const uint32_t *ids, uint32_t n) {
    const size_t leaf_start = n - 1;
#pragma omp parallel for schedule(static)
    for (int64_t g = 0; g < (int64_t)n; g++) {
        nodes[leaf_start + g].data[0] = ids[g];
        nodes[leaf_start + g].right_edge = (uint32_t)g;
    }
}

/* collision/collision.cl:65-77 (delta): common-prefix length of codes i and
 * j = i +/- offset; ties broken on sorted position (32 + clz(i ^ j)); -1 out
 * of range.  Returned as int (the reference's `char` holds -1..64 exactly). */
static inline int delta(const uint32_t *codes, uint32_t n, uint32_t i, int forward,
                        uint32_t offset) {
    if (forward && (i + offset) >= n) return -1;
    if (!forward && offset > i) return -1;
    const uint32_t j = forward ? i + offset : i - offset;
    if (codes[i] != codes[j]) return __builtin_clz(codes[i] ^ codes[j]);
    return 32 + __builtin_clz(i ^ j);
}

/* collision/collision.cl:81-121 (generateBVH), Karras 2012. */
ORC_API void orc_generate_bvh(const uint32_t *codes, orc_node *nodes, uint32_t n) {
    if (n < 2) return;
    const uint32_t leaf_start = n - 1;
#pragma omp parallel for schedule(static)
    for (int64_t g = 0; g < (int64_t)n - 1; g++) {
        const uint32_t i = (uint32_t)g;
        const int forward = delta(codes, n, i, 1, 1) > delta(codes, n, i, 0, 1);
        const int delta_min = delta(codes, n, i, !forward, 1);

        uint32_t len_max = 2;
        while (delta(codes, n, i, forward, len_max) > delta_min) len_max *= 2;

        uint32_t len = 0;
        for (uint32_t t = len_max / 2; t > 0; t /= 2)
            if (delta(codes, n, i, forward, len + t) > delta_min) len += t;

        const uint32_t j = forward ? i + len : i - len;
        const int delta_node = (unsigned char)delta(codes, n, i, forward, len);
        uint32_t s = 0;
        {
            uint32_t t = len;
            do {
                t = (t + 1) / 2;
                if (delta(codes, n, i, forward, s + t) > delta_node) s += t;
            } while (t > 1);
        }
        const uint32_t gamma = forward ? i + s : i - s - 1;
        const uint32_t lo = i < j ? i : j, hi = i < j ? j : i;
        const uint32_t child_a = (lo == gamma) ? leaf_start + gamma : gamma;
        const uint32_t child_b = (hi == gamma + 1) ? leaf_start + gamma + 1 : gamma + 1;

        nodes[i].data[0] = child_a;
        nodes[i].data[1] = child_b;
        nodes[i].right_edge = hi;
        nodes[child_a].parent = i;
        nodes[child_b].parent = i;
    }
}

/* ------------------------------------------------------------------- refit
 * Bound = {min[4], max[4]} (collision/collision.cl:123-126; 3-vector = 4
 * lanes).  collision/collision.cl:128-141 (leafBounds). */
#define DEF_LEAF_BOUNDS(NAME, T)                                                 \
    ORC_API void NAME(T *bounds, const T *coords, const T *radii,                \
                      const orc_node *nodes, uint32_t n) {                       \
        const size_t leaf_start = n - 1;                                         \
        _Pragma("omp parallel for schedule(static)")                             \
        for (int64_t g = 0; g < (int64_t)n; g++) {                               \
            size_t node = leaf_start + g;                                        \
            uint32_t c = nodes[node].data[0];                                    \
            for (int k = 0; k < 3; k++) {                                        \
                bounds[node * 8 + k] = coords[(size_t)c * 4 + k] - radii[c];     \
                bounds[node * 8 + 4 + k] = coords[(size_t)c * 4 + k] + radii[c]; \
            }                                                                    \
        }                                                                        \
    }
DEF_LEAF_BOUNDS(orc_leaf_bounds_f32, float)
DEF_LEAF_BOUNDS(orc_leaf_bounds_f64, double)

/* collision/collision.cl:143-162 (internalBounds): every leaf walks up; the
 * first arrival at a node stops, the second merges the two child boxes.
 * flags must be zero on entry (collision/collision.py:147-150). */
#define DEF_INTERNAL_BOUNDS(NAME, T, FMIN, FMAX)                                   \
    ORC_API void NAME(T *bounds, uint32_t *flags, const orc_node *nodes,           \
                      uint32_t n) {                                                \
        if (n < 2) return;                                                         \
        const size_t leaf_start = n - 1;                                           \
        _Pragma("omp parallel for schedule(static)")                               \
        for (int64_t g = 0; g < (int64_t)n; g++) {                                 \
            size_t node = leaf_start + g;                                          \
            do {                                                                   \
                node = nodes[node].parent;                                         \
                if (__atomic_fetch_add(&flags[node], 1u, __ATOMIC_ACQ_REL) < 1)    \
                    break;                                                         \
                const uint32_t *ch = nodes[node].data;                             \
                for (int k = 0; k < 3; k++) {                                      \
                    bounds[node * 8 + k] = FMIN(bounds[(size_t)ch[0] * 8 + k],     \
                                                bounds[(size_t)ch[1] * 8 + k]);    \
                    bounds[node * 8 + 4 + k] =                                     \
                        FMAX(bounds[(size_t)ch[0] * 8 + 4 + k],                    \
                             bounds[(size_t)ch[1] * 8 + 4 + k]);                   \
                }                                                                  \
            } while (node != 0);                                                   \
        }                                                                          \
    }
DEF_INTERNAL_BOUNDS(orc_internal_bounds_f32, float, fminf, fmaxf)
DEF_INTERNAL_BOUNDS(orc_internal_bounds_f64, double, fmin, fmax)

/* ---------------------------------------------------------------- traverse
 * collision/collision.cl:164-166 (checkOverlap): strict, per axis, lanes 0-2. */
#define DEF_OVERLAP(NAME, T)                                                  \
    static inline int NAME(const T *a, const T *b) {                          \
        return (a[4] > b[0]) & (a[0] < b[4]) & (a[5] > b[1]) & (a[1] < b[5]) & \
               (a[6] > b[2]) & (a[2] < b[6]);                                 \
    }
DEF_OVERLAP(overlap_f32, float)
DEF_OVERLAP(overlap_f64, double)

/* One query of collision/collision.cl:174-226 (traverse): DFS from the root
 * with a 64-entry stack seeded with the UINT_MAX sentinel; child visited iff
 * boxes overlap and child.right_edge > query_idx; leaf hits emitted as
 * (id[query], id[leaf]) child a before child b.  `emit` receives hits in
 * that order; returns the number of hits. */
#define DEF_TRAVERSE_ONE(NAME, T, OVERLAP)                                         \
    static inline uint64_t NAME(const orc_node *nodes, const T *bounds, uint32_t n, \
                                uint32_t query_idx, uint32_t *out, uint64_t room) { \
        const size_t leaf_start = n - 1;                                           \
        const T *query = bounds + (leaf_start + query_idx) * 8;                    \
        const uint32_t qid = nodes[leaf_start + query_idx].data[0];                \
        uint32_t stack[64];                                                        \
        unsigned char sp = 0;                                                      \
        uint64_t hits = 0;                                                         \
        stack[sp++] = UINT32_MAX;                                                  \
        uint32_t idx = 0;                                                          \
        do {                                                                       \
            const uint32_t ca = nodes[idx].data[0], cb = nodes[idx].data[1];       \
            int oa = OVERLAP(query, bounds + (size_t)ca * 8);                      \
            int ob = OVERLAP(query, bounds + (size_t)cb * 8);                      \
            oa &= !(nodes[ca].right_edge <= query_idx);                            \
            ob &= !(nodes[cb].right_edge <= query_idx);                            \
            if (oa && ca >= leaf_start) {                                          \
                if (hits < room) {                                                 \
                    out[hits * 2 + 0] = qid;                                       \
                    out[hits * 2 + 1] = nodes[ca].data[0];                         \
                }                                                                  \
                hits++;                                                            \
            }                                                                      \
            if (ob && cb >= leaf_start) {                                          \
                if (hits < room) {                                                 \
                    out[hits * 2 + 0] = qid;                                       \
                    out[hits * 2 + 1] = nodes[cb].data[0];                         \
                }                                                                  \
                hits++;                                                            \
            }                                                                      \
            const int ta = oa && ca < leaf_start, tb = ob && cb < leaf_start;      \
            if (!ta && !tb)                                                        \
                idx = stack[--sp];                                                 \
            else {                                                                 \
                idx = ta ? ca : cb;                                                \
                if (ta && tb) stack[sp++] = cb;                                    \
            }                                                                      \
        } while (idx != UINT32_MAX);                                               \
        return hits;                                                               \
    }
DEF_TRAVERSE_ONE(traverse_one_f32, float, overlap_f32)
DEF_TRAVERSE_ONE(traverse_one_f64, double, overlap_f64)

/* Whole kernel in work-item order q = q_begin .. q_end-1 (sequential, so the
 * pair order is deterministic: query-major, DFS order).  The first
 * `capacity` pairs are stored; the return value is the TRUE total, as
 * *next is in the reference (collision.cl:202-204).  collisions may be NULL
 * when capacity == 0 (count only, tests/test_collision_py.py:289).
 * counts (optional, n entries, indexed by q) receives per-query hit counts. */
#define DEF_TRAVERSE(NAME, T, ONE)                                                  \
    ORC_API uint64_t NAME(uint32_t *collisions, uint64_t capacity,                  \
                          const orc_node *nodes, const T *bounds, uint32_t n,       \
                          uint32_t q_begin, uint32_t q_end, uint32_t *counts) {     \
        uint64_t total = 0;                                                         \
        if (n < 2) return 0;                                                        \
        for (uint32_t q = q_begin; q < q_end; q++) {                                \
            uint64_t room = total < capacity ? capacity - total : 0;                \
            uint64_t h = ONE(nodes, bounds, n, q,                                   \
                             room ? collisions + total * 2 : NULL, room);           \
            if (counts) counts[q] = (uint32_t)h;                                    \
            total += h;                                                             \
        }                                                                           \
        return total;                                                               \
    }
DEF_TRAVERSE(orc_traverse_f32, float, traverse_one_f32)
DEF_TRAVERSE(orc_traverse_f64, double, traverse_one_f64)

/* Multi-threaded count -> scan -> write version producing exactly the same
 * list as orc_traverse_f32 (timed baseline only).  counts: n scratch u32. */
ORC_API uint64_t orc_traverse_f32_mt(uint32_t *collisions, uint64_t capacity,
                                     const orc_node *nodes, const float *bounds,
                                     uint32_t n, uint64_t *offsets) {
    if (n < 2) return 0;
#pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t q = 0; q < (int64_t)n; q++)
        offsets[q] = traverse_one_f32(nodes, bounds, n, (uint32_t)q, NULL, 0);
    uint64_t run = 0;
    for (uint32_t q = 0; q < n; q++) {
        uint64_t c = offsets[q];
        offsets[q] = run;
        run += c;
    }
    if (capacity) {
#pragma omp parallel for schedule(dynamic, 1024)
        for (int64_t q = 0; q < (int64_t)n; q++) {
            uint64_t off = offsets[q];
            if (off < capacity)
                traverse_one_f32(nodes, bounds, n, (uint32_t)q, collisions + off * 2,
                                 capacity - off);
        }
    }
    return run;
}

/* The reference's own append protocol (collision.cl:202-215): ONE walk per query,
 * rows appended through a shared counter, so their order is arbitrary -- exactly
 * what upstream produces ("Need to sort, order is undefined",
 * tests/test_collision_py.py:148).  Every thread batches its rows and reserves a
 * block of the list with one atomic add per batch (kinder to the CPU than one
 * atomic per row).  Timed baseline only; returns the TRUE total. */
ORC_API uint64_t orc_traverse_f32_single(uint32_t *collisions, uint64_t capacity,
                                         const orc_node *nodes, const float *bounds,
                                         uint32_t n, uint32_t q_begin, uint32_t q_end) {
    if (n < 2 || q_end <= q_begin) return 0;
    uint64_t next = 0;
    enum { BATCH = 4096, CHUNK = 256 };
#pragma omp parallel
    {
        uint32_t *rows = (uint32_t *)malloc(sizeof(uint32_t) * 2 * BATCH);
        uint64_t held = 0;
#pragma omp for schedule(dynamic, 1) nowait
        for (int64_t c = (int64_t)q_begin; c < (int64_t)q_end; c += CHUNK) {
            const uint32_t stop = (uint32_t)(c + CHUNK < (int64_t)q_end ? c + CHUNK : q_end);
            for (uint32_t q = (uint32_t)c; q < stop; q++) {
                uint64_t h = traverse_one_f32(nodes, bounds, n, q, rows + held * 2, BATCH - held);
                if (h > BATCH - held) {
                    /* the batch could not take this query: flush, then the query goes straight to the list */
                    uint64_t at = __atomic_fetch_add(&next, held + h, __ATOMIC_RELAXED);
                    for (uint64_t k = 0; k < held && at + k < capacity; k++) {
                        collisions[(at + k) * 2] = rows[k * 2];
                        collisions[(at + k) * 2 + 1] = rows[k * 2 + 1];
                    }
                    at += held;
                    held = 0;
                    if (at < capacity) traverse_one_f32(nodes, bounds, n, q, collisions + at * 2, capacity - at);
                } else {
                    held += h;
                }
            }
            if (held >= BATCH / 2) {
                uint64_t at = __atomic_fetch_add(&next, held, __ATOMIC_RELAXED);
                for (uint64_t k = 0; k < held && at + k < capacity; k++) {
                    collisions[(at + k) * 2] = rows[k * 2];
                    collisions[(at + k) * 2 + 1] = rows[k * 2 + 1];
                }
                held = 0;
            }
        }
        if (held) {
            uint64_t at = __atomic_fetch_add(&next, held, __ATOMIC_RELAXED);
            for (uint64_t k = 0; k < held && at + k < capacity; k++) {
                collisions[(at + k) * 2] = rows[k * 2];
                collisions[(at + k) * 2 + 1] = rows[k * 2 + 1];
            }
        }
        free(rows);
    }
    return next;
}

/* tests/test_collision_py.py:30-37 (find_collisions): O(N^2) box-overlap
 * brute force, independent of the tree.  Emits (i, j), i < j, row-major;
 * returns the total.  coords 4-lane slots. */
#define DEF_BRUTE(NAME, T)                                                       \
    ORC_API uint64_t NAME(uint32_t *pairs, uint64_t capacity, const T *coords,   \
                          const T *radii, uint32_t n) {                          \
        uint64_t total = 0;                                                      \
        for (uint32_t i = 0; i < n; i++)                                         \
            for (uint32_t j = i + 1; j < n; j++) {                               \
                int hit = 1;                                                     \
                for (int k = 0; k < 3; k++) {                                    \
                    T ilo = coords[(size_t)i * 4 + k] - radii[i];                \
                    T ihi = coords[(size_t)i * 4 + k] + radii[i];                \
                    T jlo = coords[(size_t)j * 4 + k] - radii[j];                \
                    T jhi = coords[(size_t)j * 4 + k] + radii[j];                \
                    hit &= (ilo < jhi) & (jlo < ihi);                            \
                }                                                                \
                if (hit) {                                                       \
                    if (total < capacity) {                                      \
                        pairs[total * 2] = i;                                    \
                        pairs[total * 2 + 1] = j;                                \
                    }                                                            \
                    total++;                                                     \
                }                                                                \
            }                                                                    \
        return total;                                                            \
    }
DEF_BRUTE(orc_brute_force_f32, float)
DEF_BRUTE(orc_brute_force_f64, double)

/* ------------------------------------------------------------ find_offsets
 * collision/offset.py:40-49 + collision/offset.cl:3-12: offsets[:] = n_values,
 * then for each adjacent sorted pair (a, b] <- g + 1, and offsets[g] = 0 for
 * g <= values[0].  Values widened to u64; offsets written as off_bytes wide. */
static inline void store_uint(void *p, uint64_t i, uint32_t bytes, uint64_t v) {
    if (bytes == 4) ((uint32_t *)p)[i] = (uint32_t)v;
    else ((uint64_t *)p)[i] = v;
}

ORC_API void orc_find_offsets(const void *values, uint64_t n_values, uint32_t val_bytes,
                              void *offsets, uint64_t n_offsets, uint32_t off_bytes) {
    for (uint64_t i = 0; i < n_offsets; i++) store_uint(offsets, i, off_bytes, n_values);
    if (n_values == 0) return;
    const uint64_t v0 = load_key(values, 0, val_bytes);
    for (uint64_t g = 0; g + 1 < n_values; g++) {
        const uint64_t a = load_key(values, g, val_bytes);
        const uint64_t b = load_key(values, g + 1, val_bytes);
        for (uint64_t i = a + 1; i <= b; i++) store_uint(offsets, i, off_bytes, g + 1);
        if (g <= v0) store_uint(offsets, g, off_bytes, 0);
    }
}

/* ---------------------------------------------------------- gather/scatter
 * collision/index.cl:1-13. */
ORC_API void orc_gather(const void *in, const void *idx, void *out, uint64_t n,
                        uint32_t value_bytes, uint32_t index_bytes) {
    for (uint64_t i = 0; i < n; i++)
        memcpy((char *)out + i * value_bytes,
               (const char *)in + load_key(idx, i, index_bytes) * value_bytes, value_bytes);
}

ORC_API void orc_scatter(const void *in, const void *idx, void *out, uint64_t n,
                         uint32_t value_bytes, uint32_t index_bytes) {
    for (uint64_t i = 0; i < n; i++)
        memcpy((char *)out + load_key(idx, i, index_bytes) * value_bytes,
               (const char *)in + i * value_bytes, value_bytes);
}

/* ---------------------------------------------------------------- pipeline
 * collision/collision.py:130-198 (Collider.get_collisions) end to end on the
 * host cores, float32: bounds -> codes -> stable sort -> leaves -> tree ->
 * leaf boxes -> refit -> traverse.  Scratch is allocated by the caller:
 *   codes[2n] ids[2n] nodes[2n-1] bounds[(2n-1)*8] flags[2n-1] offsets[n]
 * Returns the total pair count; stores at most `capacity` pairs in the
 * deterministic (query-major, DFS) order. */
static void pipeline_front_f32(const float *coords, const float *radii, uint32_t n,
                               uint32_t *codes, uint32_t *ids, orc_node *nodes,
                               float *bounds, uint32_t *flags) {
    float range[8];
    orc_bounds_f32x4_mt(coords, n, range);
    orc_morton_f32(codes, coords, range, n);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) ids[i] = (uint32_t)i;
    orc_sort_pairs(codes, codes + n, ids, ids + n, n, 4, 4);
    memset(flags, 0, sizeof(uint32_t) * (2 * (size_t)n - 1));
    orc_fill_leaves(nodes, ids + n, n);
    orc_generate_bvh(codes + n, nodes, n);
    orc_leaf_bounds_f32(bounds, coords, radii, nodes, n);
    orc_internal_bounds_f32(bounds, flags, nodes, n);
}

ORC_API uint64_t orc_pipeline_f32(const float *coords, const float *radii, uint32_t n,
                                  uint32_t *collisions, uint64_t capacity,
                                  uint32_t *codes, uint32_t *ids, orc_node *nodes,
                                  float *bounds, uint32_t *flags, uint64_t *offsets) {
    pipeline_front_f32(coords, radii, n, codes, ids, nodes, bounds, flags);
    return orc_traverse_f32_mt(collisions, capacity, nodes, bounds, n, offsets);
}

/* The same pipeline with the reference's single walk + counter append (rows in
 * arbitrary order, as upstream): what bench.py times as the CPU baseline. */
ORC_API uint64_t orc_pipeline_f32_single(const float *coords, const float *radii, uint32_t n,
                                         uint32_t *collisions, uint64_t capacity,
                                         uint32_t *codes, uint32_t *ids, orc_node *nodes,
                                         float *bounds, uint32_t *flags) {
    pipeline_front_f32(coords, radii, n, codes, ids, nodes, bounds, flags);
    return orc_traverse_f32_single(collisions, capacity, nodes, bounds, n, 0, n);
}
