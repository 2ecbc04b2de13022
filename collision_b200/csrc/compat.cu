// compat.cu -- the reference's KERNEL-LEVEL contracts for the non-collision programs.
//
// Upstream's tests launch block_sort / scatter (collision/radix.cl:48-139), local_scan /
// block_scan (collision/scan.cl:5-36), find_offsets (collision/offset.cl:3-12) and
// bounds1 / bounds2 (collision/reduce.cl:5-58) directly, with their own launch geometry
// (work-groups of `local_size` items covering 2 * local_size elements, digit-major per-group
// histograms, per-group partials), so that geometry is part of the public surface
// (SURVEY.md section 1, L0 row).  The hot path never runs these: RadixSorter.sort,
// PrefixScanner.prefix_sum, Bounds.reduce and OffsetFinder.find_offsets go through sort.cu,
// scan.cu, reduce.cu and index.cu.  These kernels exist so that
// `program.kernels[name](cq, global_size, local_size, *args)` keeps working with the
// reference's argument lists and produces the reference's buffers; they are written for
// clarity, one CUDA block per OpenCL work-group, and are not tuned.
//
// bounds1 / bounds2 are also the generic accumulator-list Reducer (collision/reduce.py:9-22):
// any list of (initialiser, min | max | ADD) pairs, any ngroups x group_size geometry --
// including upstream's quirk that a non-power-of-two ngroups silently drops the tail
// partials in bounds2 (reduce.cl:42-52), reproduced here literally.
#include "common.cuh"

namespace {

constexpr int COMPAT_THREADS = 256;

__device__ __forceinline__ uint64_t load_uint(const void* p, uint64_t i, uint32_t bytes) {
    return bytes == 4 ? (uint64_t)reinterpret_cast<const uint32_t*>(p)[i] : reinterpret_cast<const uint64_t*>(p)[i];
}
__device__ __forceinline__ void store_uint(void* p, uint64_t i, uint32_t bytes, uint64_t v) {
    if (bytes == 4) reinterpret_cast<uint32_t*>(p)[i] = (uint32_t)v;
    else reinterpret_cast<uint64_t*>(p)[i] = v;
}

// ---- local_scan (scan.cl:5-30): exclusive scan of each group of `items` values in place;
// the group's total goes to block_sums[group] when block_sums is given.
__global__ void __launch_bounds__(COMPAT_THREADS)
local_scan_kernel(uint32_t* __restrict__ data, uint32_t* __restrict__ block_sums, uint32_t items) {
    __shared__ uint32_t part[COMPAT_THREADS];
    uint32_t* d = data + (size_t)blockIdx.x * items;
    const uint32_t per = (items + COMPAT_THREADS - 1) / COMPAT_THREADS;
    const uint32_t lo = min(threadIdx.x * per, items), hi = min(lo + per, items);
    uint32_t sum = 0;
    for (uint32_t i = lo; i < hi; i++) sum += d[i];
    part[threadIdx.x] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (int t = 0; t < COMPAT_THREADS; t++) {
            const uint32_t c = part[t];
            part[t] = run;
            run += c;
        }
        if (block_sums) block_sums[blockIdx.x] = run;
    }
    __syncthreads();
    uint32_t run = part[threadIdx.x];
    for (uint32_t i = lo; i < hi; i++) {
        const uint32_t c = d[i];
        d[i] = run;
        run += c;
    }
}

// ---- block_scan (scan.cl:32-36): data[i] += block_sums[group of i]
__global__ void __launch_bounds__(COMPAT_THREADS)
block_scan_kernel(uint32_t* __restrict__ data, const uint32_t* __restrict__ block_sums, uint64_t n, uint32_t items) {
    const uint64_t i = (uint64_t)blockIdx.x * COMPAT_THREADS + threadIdx.x;
    if (i < n) data[i] += block_sums[i / items];
}

// ---- block_sort (radix.cl:48-102): every group stable-sorts its `items` keys (and values)
// by the digit, in place, and writes its digit counts to histogram[digit * n_groups + group].
// smem: [dst: items u32][hist: 2^bits u32][keys: items * key_bytes][values: items * value_bytes]
__global__ void __launch_bounds__(COMPAT_THREADS)
block_sort_kernel(void* __restrict__ keys, void* __restrict__ values, uint32_t* __restrict__ histogram,
                  uint32_t items, uint32_t key_bytes, uint32_t value_bytes, uint32_t bits, uint32_t pass) {
    extern __shared__ __align__(16) unsigned char smem[];
    uint32_t* dst = reinterpret_cast<uint32_t*>(smem);
    uint32_t* hist = dst + items;
    const uint32_t ndig = 1u << bits;
    unsigned char* skeys = reinterpret_cast<unsigned char*>(hist + ndig);
    skeys += (16 - ((size_t)skeys & 15)) & 15;
    unsigned char* svals = skeys + (((size_t)items * key_bytes + 15) & ~(size_t)15);
    const size_t start = (size_t)blockIdx.x * items;
    const uint32_t shift = pass * bits;
    const uint64_t mask = ndig - 1;
    unsigned char* gk = reinterpret_cast<unsigned char*>(keys) + start * key_bytes;
    unsigned char* gv = values ? reinterpret_cast<unsigned char*>(values) + start * value_bytes : nullptr;
    for (uint32_t i = threadIdx.x; i < items * key_bytes / 4; i += COMPAT_THREADS)
        reinterpret_cast<uint32_t*>(skeys)[i] = reinterpret_cast<uint32_t*>(gk)[i];
    if (gv)
        for (uint32_t i = threadIdx.x; i < items * value_bytes / 4; i += COMPAT_THREADS)
            reinterpret_cast<uint32_t*>(svals)[i] = reinterpret_cast<uint32_t*>(gv)[i];
    for (uint32_t d = threadIdx.x; d < ndig; d += COMPAT_THREADS) hist[d] = 0;
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < items; i += COMPAT_THREADS)
        atomicAdd(&hist[(load_uint(skeys, i, key_bytes) >> shift) & mask], 1u);
    __syncthreads();
    for (uint32_t d = threadIdx.x; d < ndig; d += COMPAT_THREADS)
        histogram[(size_t)d * gridDim.x + blockIdx.x] = hist[d];
    __syncthreads();
    if (threadIdx.x == 0) {   // counts -> first slot of each digit
        uint32_t run = 0;
        for (uint32_t d = 0; d < ndig; d++) {
            const uint32_t c = hist[d];
            hist[d] = run;
            run += c;
        }
    }
    __syncthreads();
    // stable placement: one thread per digit walks the group in order
    for (uint32_t d = threadIdx.x; d < ndig; d += COMPAT_THREADS) {
        uint32_t at = hist[d];
        for (uint32_t i = 0; i < items; i++)
            if (((load_uint(skeys, i, key_bytes) >> shift) & mask) == d) dst[i] = at++;
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < items; i += COMPAT_THREADS) {
        const uint32_t to = dst[i];
        for (uint32_t w = 0; w < key_bytes / 4; w++)
            reinterpret_cast<uint32_t*>(gk)[(size_t)to * (key_bytes / 4) + w] =
                reinterpret_cast<uint32_t*>(skeys)[(size_t)i * (key_bytes / 4) + w];
        if (gv)
            for (uint32_t w = 0; w < value_bytes / 4; w++)
                reinterpret_cast<uint32_t*>(gv)[(size_t)to * (value_bytes / 4) + w] =
                    reinterpret_cast<uint32_t*>(svals)[(size_t)i * (value_bytes / 4) + w];
    }
}

// ---- scatter (radix.cl:104-139): group g moves its block-sorted items to
// out[offset[d * n_groups + g] + i - (first slot of digit d inside the group)]
__global__ void __launch_bounds__(COMPAT_THREADS)
scatter_kernel(const void* __restrict__ keys, void* __restrict__ out_keys, const void* __restrict__ values,
               void* __restrict__ out_values, const uint32_t* __restrict__ offset,
               const uint32_t* __restrict__ histogram, uint32_t items, uint32_t key_bytes, uint32_t value_bytes,
               uint32_t bits, uint32_t pass) {
    extern __shared__ __align__(16) unsigned char smem[];
    uint32_t* local_offset = reinterpret_cast<uint32_t*>(smem);
    const uint32_t ndig = 1u << bits;
    uint32_t* local_start = local_offset + ndig;
    for (uint32_t d = threadIdx.x; d < ndig; d += COMPAT_THREADS) {
        local_offset[d] = offset[(size_t)d * gridDim.x + blockIdx.x];
        local_start[d] = histogram[(size_t)d * gridDim.x + blockIdx.x];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (uint32_t d = 0; d < ndig; d++) {
            const uint32_t c = local_start[d];
            local_start[d] = run;
            run += c;
        }
    }
    __syncthreads();
    const size_t start = (size_t)blockIdx.x * items;
    const uint32_t shift = pass * bits;
    const uint64_t mask = ndig - 1;
    for (uint32_t i = threadIdx.x; i < items; i += COMPAT_THREADS) {
        const uint64_t k = load_uint(keys, start + i, key_bytes);
        const uint32_t d = (uint32_t)((k >> shift) & mask);
        const size_t to = (size_t)local_offset[d] + i - local_start[d];
        store_uint(out_keys, to, key_bytes, k);
        if (values && out_values)
            for (uint32_t w = 0; w < value_bytes / 4; w++)
                reinterpret_cast<uint32_t*>(out_values)[to * (value_bytes / 4) + w] =
                    reinterpret_cast<const uint32_t*>(values)[(start + i) * (value_bytes / 4) + w];
    }
}

// ---- find_offsets (offset.cl:3-12), literally: no fill (that is the host's, offset.py:43-46)
__global__ void __launch_bounds__(COMPAT_THREADS)
find_offsets_literal_kernel(const void* __restrict__ values, uint64_t n_threads, uint32_t value_bytes,
                            void* __restrict__ offsets, uint32_t offset_bytes) {
    const uint64_t g = (uint64_t)blockIdx.x * COMPAT_THREADS + threadIdx.x;
    if (g >= n_threads) return;
    const uint64_t a = load_uint(values, g, value_bytes), b = load_uint(values, g + 1, value_bytes);
    for (uint64_t i = a + 1; i <= b; i++) store_uint(offsets, i, offset_bytes, g + 1);
    if (g <= load_uint(values, 0, value_bytes)) store_uint(offsets, g, offset_bytes, 0);
}

// ---- bounds1 / bounds2 (reduce.cl:5-58) for any accumulator list
constexpr int MAX_ACC = 8, MAX_LANES = 16;
enum { OP_MIN = 0, OP_MAX = 1, OP_ADD = 2 };
struct AccSpec {
    int n_acc;
    int op[MAX_ACC];
    double init[MAX_ACC];
};

template <typename T> __device__ __forceinline__ T apply_op(int op, T a, T b) {
    // min / max as OpenCL's (fmin / fmax semantics for floats: a NaN operand loses), ADD = a + b
    if (op == OP_ADD) return a + b;
    if (op == OP_MIN) return b < a || a != a ? b : a;
    return b > a || a != a ? b : a;
}

// values: n slots of `lanes` scalars; group_accs[group][acc][lane]
template <typename T>
__global__ void reduce1_kernel(const T* __restrict__ values, uint64_t n, uint32_t lanes, AccSpec spec,
                               T* __restrict__ group_accs) {
    extern __shared__ __align__(16) unsigned char smem[];
    T* scratch = reinterpret_cast<T*>(smem);                  // [local_id][acc][lane]
    const uint32_t stride = spec.n_acc * lanes;
    T acc[MAX_ACC * MAX_LANES];
    for (int a = 0; a < spec.n_acc; a++)
        for (uint32_t l = 0; l < lanes; l++) acc[a * lanes + l] = (T)spec.init[a];
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        for (int a = 0; a < spec.n_acc; a++)
            for (uint32_t l = 0; l < lanes; l++)
                acc[a * lanes + l] = apply_op(spec.op[a], acc[a * lanes + l], values[i * lanes + l]);
    for (uint32_t k = 0; k < stride; k++) scratch[(size_t)threadIdx.x * stride + k] = acc[k];
    __syncthreads();
    for (uint32_t o = blockDim.x / 2; o > 0; o /= 2) {        // (reduce.cl:26-33: drops the tail if not a power of two)
        if (threadIdx.x < o)
            for (int a = 0; a < spec.n_acc; a++)
                for (uint32_t l = 0; l < lanes; l++) {
                    const size_t k = a * lanes + l;
                    scratch[(size_t)threadIdx.x * stride + k] = apply_op(
                        spec.op[a], scratch[(size_t)threadIdx.x * stride + k], scratch[(size_t)(threadIdx.x + o) * stride + k]);
                }
        __syncthreads();
    }
    if (threadIdx.x == 0)
        for (uint32_t k = 0; k < stride; k++) group_accs[(size_t)blockIdx.x * stride + k] = scratch[k];
}

// one block of `ngroups` threads; the partials are reduced IN PLACE in global memory as upstream
template <typename T>
__global__ void reduce2_kernel(T* group_accs, T* __restrict__ output, uint32_t lanes, AccSpec spec) {
    const uint32_t stride = spec.n_acc * lanes;
    for (uint32_t o = blockDim.x / 2; o > 0; o /= 2) {        // reduce.cl:42-52
        if (threadIdx.x < o)
            for (int a = 0; a < spec.n_acc; a++)
                for (uint32_t l = 0; l < lanes; l++) {
                    const size_t k = a * lanes + l;
                    group_accs[(size_t)threadIdx.x * stride + k] = apply_op(
                        spec.op[a], group_accs[(size_t)threadIdx.x * stride + k],
                        group_accs[(size_t)(threadIdx.x + o) * stride + k]);
                }
        __syncthreads();
    }
    if (threadIdx.x == 0)
        for (uint32_t k = 0; k < stride; k++) output[k] = group_accs[k];
}

inline size_t scalar_bytes(int dtype) {
    switch (dtype) {
        case COL_F32: case COL_U32: case COL_I32: return 4;
        case COL_F64: case COL_U64: case COL_I64: return 8;
    }
    return 0;
}

inline int make_spec(AccSpec& spec, const int* h_ops, const double* h_inits, uint32_t n_acc) {
    if (n_acc == 0 || n_acc > MAX_ACC || !h_ops) return COL_ERR_BAD_ARG;
    spec.n_acc = (int)n_acc;
    for (uint32_t a = 0; a < n_acc; a++) {
        if (h_ops[a] < OP_MIN || h_ops[a] > OP_ADD) return COL_ERR_BAD_ARG;
        spec.op[a] = h_ops[a];
        spec.init[a] = h_inits ? h_inits[a] : 0.0;
    }
    return 0;
}

}  // namespace

COL_API int col_compat_local_scan(uint32_t* data, uint32_t* block_sums, uint64_t n_groups, uint32_t items_per_group,
                                  void* stream) {
    if (!data || items_per_group == 0 || n_groups > 0x7fffffffu) return COL_ERR_BAD_ARG;
    if (n_groups == 0) return 0;
    local_scan_kernel<<<(unsigned)n_groups, COMPAT_THREADS, 0, col_stream(stream)>>>(data, block_sums, items_per_group);
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_compat_block_scan(uint32_t* data, const uint32_t* block_sums, uint64_t n_items,
                                  uint32_t items_per_group, void* stream) {
    if (!data || !block_sums || items_per_group == 0) return COL_ERR_BAD_ARG;
    if (n_items == 0) return 0;
    block_scan_kernel<<<col_blocks(n_items, COMPAT_THREADS), COMPAT_THREADS, 0, col_stream(stream)>>>(
        data, block_sums, n_items, items_per_group);
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_compat_block_sort(void* keys, void* values, uint32_t* histogram, uint64_t n_groups,
                                  uint32_t items_per_group, uint32_t key_bytes, uint32_t value_bytes,
                                  uint32_t radix_bits, uint32_t pass, void* stream) {
    if (!keys || !histogram || (key_bytes != 4 && key_bytes != 8) || (value_bytes & 3) || radix_bits == 0 ||
        radix_bits > 12 || (pass + 1) * radix_bits > key_bytes * 8 || items_per_group == 0 || n_groups > 0x7fffffffu)
        return COL_ERR_BAD_ARG;
    if (n_groups == 0) return 0;
    const size_t smem = sizeof(uint32_t) * ((size_t)items_per_group + ((size_t)1 << radix_bits)) + 32 +
                        (((size_t)items_per_group * key_bytes + 15) & ~(size_t)15) +
                        (values ? (size_t)items_per_group * value_bytes : 0);
    if (smem > 200 * 1024) return COL_ERR_TOO_LARGE;
    COL_CUDA(cudaFuncSetAttribute(block_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    block_sort_kernel<<<(unsigned)n_groups, COMPAT_THREADS, smem, col_stream(stream)>>>(
        keys, values, histogram, items_per_group, key_bytes, values ? value_bytes : 0, radix_bits, pass);
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_compat_scatter(const void* keys, void* out_keys, const void* values, void* out_values,
                               const uint32_t* offset, const uint32_t* histogram, uint64_t n_groups,
                               uint32_t items_per_group, uint32_t key_bytes, uint32_t value_bytes, uint32_t radix_bits,
                               uint32_t pass, void* stream) {
    if (!keys || !out_keys || !offset || !histogram || (key_bytes != 4 && key_bytes != 8) || (value_bytes & 3) ||
        radix_bits == 0 || radix_bits > 12 || (pass + 1) * radix_bits > key_bytes * 8 || items_per_group == 0 ||
        n_groups > 0x7fffffffu)
        return COL_ERR_BAD_ARG;
    if (n_groups == 0) return 0;
    const size_t smem = sizeof(uint32_t) * 2 * ((size_t)1 << radix_bits);
    scatter_kernel<<<(unsigned)n_groups, COMPAT_THREADS, smem, col_stream(stream)>>>(
        keys, out_keys, values, out_values, offset, histogram, items_per_group, key_bytes, value_bytes, radix_bits, pass);
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_compat_find_offsets(const void* values, uint64_t n_threads, uint32_t value_bytes, void* offsets,
                                    uint32_t offset_bytes, void* stream) {
    if (!values || !offsets || (value_bytes != 4 && value_bytes != 8) || (offset_bytes != 4 && offset_bytes != 8))
        return COL_ERR_BAD_ARG;
    if (n_threads == 0) return 0;
    find_offsets_literal_kernel<<<col_blocks(n_threads, COMPAT_THREADS), COMPAT_THREADS, 0, col_stream(stream)>>>(
        values, n_threads, value_bytes, offsets, offset_bytes);
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_compat_reduce1(const void* values, uint64_t n, int dtype, uint32_t lanes, const int* h_ops,
                               const double* h_inits, uint32_t n_acc, void* group_accs, uint32_t ngroups,
                               uint32_t group_size, void* stream) {
    AccSpec spec;
    const int rc = make_spec(spec, h_ops, h_inits, n_acc);
    if (rc) return rc;
    const size_t sb = scalar_bytes(dtype);
    if (!values || !group_accs || !sb || lanes == 0 || lanes > MAX_LANES || ngroups == 0 || group_size == 0 ||
        group_size > 1024)
        return sb ? COL_ERR_BAD_ARG : COL_ERR_UNSUPPORTED;
    const size_t smem = (size_t)group_size * n_acc * lanes * sb;
    if (smem > 200 * 1024) return COL_ERR_TOO_LARGE;
    cudaStream_t s = col_stream(stream);
#define COL_R1(T)                                                                                                    \
    do {                                                                                                             \
        COL_CUDA(cudaFuncSetAttribute(reduce1_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));   \
        reduce1_kernel<T><<<ngroups, group_size, smem, s>>>((const T*)values, n, lanes, spec, (T*)group_accs);       \
    } while (0)
    switch (dtype) {
        case COL_F32: COL_R1(float); break;
        case COL_F64: COL_R1(double); break;
        case COL_U32: COL_R1(uint32_t); break;
        case COL_I32: COL_R1(int32_t); break;
        case COL_U64: COL_R1(unsigned long long); break;
        case COL_I64: COL_R1(long long); break;
    }
#undef COL_R1
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_compat_reduce2(void* group_accs, void* output, int dtype, uint32_t lanes, const int* h_ops,
                               uint32_t n_acc, uint32_t ngroups, void* stream) {
    AccSpec spec;
    const int rc = make_spec(spec, h_ops, nullptr, n_acc);
    if (rc) return rc;
    if (!group_accs || !output || !scalar_bytes(dtype) || lanes == 0 || lanes > MAX_LANES || ngroups == 0 ||
        ngroups > 1024)
        return scalar_bytes(dtype) ? COL_ERR_BAD_ARG : COL_ERR_UNSUPPORTED;
    cudaStream_t s = col_stream(stream);
    switch (dtype) {
        case COL_F32: reduce2_kernel<float><<<1, ngroups, 0, s>>>((float*)group_accs, (float*)output, lanes, spec); break;
        case COL_F64: reduce2_kernel<double><<<1, ngroups, 0, s>>>((double*)group_accs, (double*)output, lanes, spec); break;
        case COL_U32: reduce2_kernel<uint32_t><<<1, ngroups, 0, s>>>((uint32_t*)group_accs, (uint32_t*)output, lanes, spec); break;
        case COL_I32: reduce2_kernel<int32_t><<<1, ngroups, 0, s>>>((int32_t*)group_accs, (int32_t*)output, lanes, spec); break;
        case COL_U64: reduce2_kernel<unsigned long long><<<1, ngroups, 0, s>>>((unsigned long long*)group_accs, (unsigned long long*)output, lanes, spec); break;
        case COL_I64: reduce2_kernel<long long><<<1, ngroups, 0, s>>>((long long*)group_accs, (long long*)output, lanes, spec); break;
    }
    COL_CHECK_LAUNCH();
    return 0;
}
