// reduce.cu -- scene-bounds reduction (min/max) and vector sum.
//
// Replaces bounds1 + bounds2 (collision/reduce.cl:5-58; launched by
// collision/reduce.py:62-76) with ONE launch: a grid sized to the SM count
// strides over the slots with 16-byte vector loads, every thread keeps a
// LANES-wide accumulator pair in registers, lanes never mix (the pad lane of a
// 3-vector slot is carried along but can never contaminate x/y/z), warps
// combine by shuffle, blocks by shared memory, and the last block to finish
// (atomic ticket) folds the per-block partials and writes the result.
// min/max are exact and order-free, so the result is bit-identical to the
// reference for any launch geometry.  Roofline: HBM, 16 B per f32x4 slot.
#include "common.cuh"

namespace {

constexpr int RED_THREADS = 512;
constexpr int RED_BLOCKS = COL_NUM_SMS * 2;
constexpr int RED_MAX_LANES = 16;

struct OpMinMax {
    static constexpr int NACC = 2;
    template <typename T> __device__ static T init(int a) { return a == 0 ? (T)INFINITY : (T)-INFINITY; }
    __device__ static float apply(int a, float x, float y) { return a == 0 ? fminf(x, y) : fmaxf(x, y); }
    __device__ static double apply(int a, double x, double y) { return a == 0 ? fmin(x, y) : fmax(x, y); }
};
struct OpSum {
    static constexpr int NACC = 1;
    template <typename T> __device__ static T init(int) { return (T)0; }
    template <typename T> __device__ static T apply(int, T x, T y) { return x + y; }
};

// scratch layout: [0] ticket u32 (zero between calls), pad to 256 B, then
// RED_BLOCKS * NACC * RED_MAX_LANES doubles (sized for the widest case).
constexpr size_t RED_PART_OFF = 256;

template <typename T, int LANES, typename Op>
__global__ void __launch_bounds__(RED_THREADS)
reduce_kernel(const T* __restrict__ vals, uint64_t n, T* __restrict__ out, unsigned* ticket, T* partials) {
    constexpr int NACC = Op::NACC;
    constexpr int VEC = 16 / sizeof(T);                        // scalars per 16-byte load
    constexpr int SLOT_VECS = (LANES >= VEC) ? LANES / VEC : 1;  // 16-byte loads per slot
    constexpr int VEC_SLOTS = (LANES >= VEC) ? 1 : VEC / LANES;  // slots per 16-byte load

    T acc[NACC][LANES];
#pragma unroll
    for (int a = 0; a < NACC; a++)
#pragma unroll
        for (int l = 0; l < LANES; l++) acc[a][l] = Op::template init<T>(a);

    const uint64_t tid = (uint64_t)blockIdx.x * RED_THREADS + threadIdx.x;
    const uint64_t nthreads = (uint64_t)gridDim.x * RED_THREADS;
    const bool aligned = (reinterpret_cast<uintptr_t>(vals) & 15u) == 0;

    // units of work: groups of VEC_SLOTS slots = SLOT_VECS vector loads
    const uint64_t n_units = aligned ? n / VEC_SLOTS : 0;
    const uint4* v4 = reinterpret_cast<const uint4*>(vals);
    constexpr int UNROLL = (SLOT_VECS >= 4) ? 1 : 4 / SLOT_VECS;  // >= 4 independent 16-byte loads in flight
    uint64_t u = tid;
    for (; u + (UNROLL - 1) * nthreads < n_units; u += UNROLL * nthreads) {
        T x[UNROLL][SLOT_VECS * VEC];
#pragma unroll
        for (int r = 0; r < UNROLL; r++)
#pragma unroll
            for (int k = 0; k < SLOT_VECS; k++)
                *reinterpret_cast<uint4*>(&x[r][k * VEC]) = ld_stream_u4(v4 + (u + r * nthreads) * SLOT_VECS + k);
#pragma unroll
        for (int r = 0; r < UNROLL; r++)
#pragma unroll
            for (int e = 0; e < SLOT_VECS * VEC; e++)
#pragma unroll
                for (int a = 0; a < NACC; a++) acc[a][e % LANES] = Op::apply(a, acc[a][e % LANES], x[r][e]);
    }
    for (; u < n_units; u += nthreads) {
        T x[SLOT_VECS * VEC];
#pragma unroll
        for (int k = 0; k < SLOT_VECS; k++)
            *reinterpret_cast<uint4*>(&x[k * VEC]) = ld_stream_u4(v4 + u * SLOT_VECS + k);
#pragma unroll
        for (int e = 0; e < SLOT_VECS * VEC; e++)
#pragma unroll
            for (int a = 0; a < NACC; a++) acc[a][e % LANES] = Op::apply(a, acc[a][e % LANES], x[e]);
    }
    // scalar tail (and the whole array when the pointer is not 16-byte aligned)
    for (uint64_t s = n_units * VEC_SLOTS + tid; s < n; s += nthreads)
#pragma unroll
        for (int l = 0; l < LANES; l++) {
            T x = vals[s * LANES + l];
#pragma unroll
            for (int a = 0; a < NACC; a++) acc[a][l] = Op::apply(a, acc[a][l], x);
        }

    // warp -> block
    __shared__ T sh[RED_THREADS / 32][NACC][LANES];
    __shared__ bool is_last;
#pragma unroll
    for (int a = 0; a < NACC; a++)
#pragma unroll
        for (int l = 0; l < LANES; l++) {
            T v = acc[a][l];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v = Op::apply(a, v, __shfl_xor_sync(0xffffffffu, v, o));
            if (lane_id() == 0) sh[threadIdx.x >> 5][a][l] = v;
        }
    __syncthreads();
    if (threadIdx.x < NACC * LANES) {
        const int a = threadIdx.x / LANES, l = threadIdx.x % LANES;
        T v = sh[0][a][l];
        for (int w = 1; w < RED_THREADS / 32; w++) v = Op::apply(a, v, sh[w][a][l]);
        partials[(size_t)blockIdx.x * NACC * LANES + threadIdx.x] = v;
        __threadfence();
    }
    __syncthreads();
    if (threadIdx.x == 0) is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    // fold the per-block partials with the whole block: thread t owns column t % W and the rows
    // t / W, t / W + ROWS, ...; the loads are independent (no serial chain of L2 round trips)
    constexpr int W = NACC * LANES, ROWS = RED_THREADS / W;
    __shared__ T fold[RED_THREADS];
    {
        const int a = (threadIdx.x % W) / LANES;
        T v = Op::template init<T>(a);
        for (unsigned b = threadIdx.x / W; b < gridDim.x; b += ROWS)
            v = Op::apply(a, v, __ldcg(partials + (size_t)b * W + threadIdx.x % W));
        fold[threadIdx.x] = v;
        __syncthreads();
        for (int s = RED_THREADS / 2; s >= W; s >>= 1) {
            if ((int)threadIdx.x < s) fold[threadIdx.x] = Op::apply(a, fold[threadIdx.x], fold[threadIdx.x + s]);
            __syncthreads();
        }
        if (threadIdx.x < W) out[threadIdx.x] = fold[threadIdx.x];
    }
    if (threadIdx.x == 0) *ticket = 0;  // ready for the next call on this scratch
}

template <typename T, typename Op>
int launch_reduce(const void* values, uint64_t n, uint32_t lanes, void* out, void* scratch, cudaStream_t s) {
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
    T* partials = reinterpret_cast<T*>(reinterpret_cast<char*>(scratch) + RED_PART_OFF);
    COL_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned), s));
    uint64_t want = (n + RED_THREADS - 1) / RED_THREADS;
    unsigned blocks = (unsigned)(want < 1 ? 1 : (want > RED_BLOCKS ? RED_BLOCKS : want));
#define LAUNCH(L)                                                                                     \
    case L:                                                                                           \
        reduce_kernel<T, L, Op><<<blocks, RED_THREADS, 0, s>>>((const T*)values, n, (T*)out, ticket, partials); \
        break;
    switch (lanes) {
        LAUNCH(1) LAUNCH(2) LAUNCH(4) LAUNCH(8) LAUNCH(16)
        default: return COL_ERR_BAD_ARG;
    }
#undef LAUNCH
    COL_CHECK_LAUNCH();
    return 0;
}

}  // namespace

COL_API size_t col_reduce_scratch_bytes(void) {
    return RED_PART_OFF + (size_t)RED_BLOCKS * 2 * RED_MAX_LANES * sizeof(double);
}

COL_API int col_bounds(const void* values, uint64_t n, int dtype, uint32_t lanes, void* out_minmax,
                       void* scratch, void* stream) {
    if (!values || !out_minmax || !scratch) return COL_ERR_BAD_ARG;
    if (dtype == COL_F32) return launch_reduce<float, OpMinMax>(values, n, lanes, out_minmax, scratch, col_stream(stream));
    if (dtype == COL_F64) return launch_reduce<double, OpMinMax>(values, n, lanes, out_minmax, scratch, col_stream(stream));
    return COL_ERR_UNSUPPORTED;
}

COL_API int col_sum(const void* values, uint64_t n, int dtype, uint32_t lanes, void* out_sum, void* scratch,
                    void* stream) {
    if (!values || !out_sum || !scratch) return COL_ERR_BAD_ARG;
    if (dtype == COL_F32) return launch_reduce<float, OpSum>(values, n, lanes, out_sum, scratch, col_stream(stream));
    if (dtype == COL_F64) return launch_reduce<double, OpSum>(values, n, lanes, out_sum, scratch, col_stream(stream));
    // integer sums wrap modulo 2^32 / 2^64 (OpenCL integer arithmetic: two's complement for the signed
    // types as well), so the unsigned instantiation serves both and the result is exact for any order
    if (dtype == COL_U32 || dtype == COL_I32)
        return launch_reduce<uint32_t, OpSum>(values, n, lanes, out_sum, scratch, col_stream(stream));
    if (dtype == COL_U64 || dtype == COL_I64)
        return launch_reduce<unsigned long long, OpSum>(values, n, lanes, out_sum, scratch, col_stream(stream));
    return COL_ERR_UNSUPPORTED;
}
