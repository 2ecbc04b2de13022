// morton.cu -- 30-bit Morton codes + iota ids + key padding in one pass.
//
// Replaces range (collision/collision.cl:8-10), the 0xFFFFFFFF key fill
// (collision/collision.py:138-142) and calculateCodes / morton / expandBits
// (collision/collision.cl:14-40).  Pure streaming: one 16-byte (f32) or
// 32-byte (f64) coalesced load per sphere, 4 + 4 bytes stored.  The arithmetic
// is the reference's, operation for operation: IEEE subtract, correctly
// rounded divide (-prec-div=true), multiply by 1023, clamp as
// fmin(fmax(x, 0), 1023) (OpenCL clamp; NaN -> 0), truncating conversion.
#include "common.cuh"

namespace {

__device__ __forceinline__ uint32_t expand_bits(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__device__ __forceinline__ uint32_t quantise(float p, float lo, float hi) {
    float q = __fdiv_rn(__fsub_rn(p, lo), __fsub_rn(hi, lo));
    q = __fmul_rn(q, 1023.0f);
    q = fminf(fmaxf(q, 0.0f), 1023.0f);
    return (uint32_t)q;  // cvt.rzi
}
__device__ __forceinline__ uint32_t quantise(double p, double lo, double hi) {
    double q = __ddiv_rn(__dsub_rn(p, lo), __dsub_rn(hi, lo));
    q = __dmul_rn(q, 1023.0);
    q = fmin(fmax(q, 0.0), 1023.0);
    return (uint32_t)q;
}

template <typename T>
__global__ void __launch_bounds__(256)
morton_kernel(uint32_t* __restrict__ codes, uint32_t* __restrict__ ids, const T* __restrict__ coords,
              const T* __restrict__ range, uint32_t n, uint32_t n_padded) {
    const T lox = range[0], loy = range[1], loz = range[2];
    const T hix = range[4], hiy = range[5], hiz = range[6];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_padded; i += gridDim.x * blockDim.x) {
        uint32_t code = 0xFFFFFFFFu;
        if (i < n) {
            T x, y, z;
            if constexpr (sizeof(T) == 4) {
                float4 c = ld_stream_f4(reinterpret_cast<const float4*>(coords) + i);
                x = c.x; y = c.y; z = c.z;
            } else {
                const double2* p = reinterpret_cast<const double2*>(coords) + (size_t)i * 2;
                double2 a = __ldg(p), b = __ldg(p + 1);
                x = a.x; y = a.y; z = b.x;
            }
            code = (expand_bits(quantise(x, lox, hix)) << 2) + (expand_bits(quantise(y, loy, hiy)) << 1) +
                   expand_bits(quantise(z, loz, hiz));
        }
        codes[i] = code;
        if (ids) ids[i] = i;
    }
}

}  // namespace

COL_API int col_morton(uint32_t* codes, uint32_t* ids, const void* coords4, const void* range2x4, int dtype,
                       uint32_t n, uint32_t n_padded, void* stream) {
    if (!codes || !coords4 || !range2x4 || n_padded < n) return COL_ERR_BAD_ARG;
    if (n_padded == 0) return 0;
    unsigned blocks = col_blocks(n_padded, 256);
    if (blocks > COL_NUM_SMS * 16) blocks = COL_NUM_SMS * 16;
    if (dtype == COL_F32)
        morton_kernel<float><<<blocks, 256, 0, col_stream(stream)>>>(codes, ids, (const float*)coords4,
                                                                     (const float*)range2x4, n, n_padded);
    else if (dtype == COL_F64)
        morton_kernel<double><<<blocks, 256, 0, col_stream(stream)>>>(codes, ids, (const double*)coords4,
                                                                      (const double*)range2x4, n, n_padded);
    else
        return COL_ERR_UNSUPPORTED;
    COL_CHECK_LAUNCH();
    return 0;
}
