// index.cu -- gather / scatter / find_offsets / fill: the small standalone
// utilities of the reference package.
//
// gather/scatter: collision/index.cl:1-13 (Indexer, collision/index.py:37-55).
// Values are moved as 4-, 8- or 16-byte words: one thread per word, so that a
// wide value (e.g. a 32-byte f64x4 slot) is copied by neighbouring lanes with
// coalesced 16-byte accesses on the contiguous side.
//
// find_offsets: collision/offset.cl:3-12 + the host fill of
// collision/offset.py:43-46.  Re-expressed as one thread per OUTPUT entry doing
// a lower-bound binary search in the sorted values: no pre-fill, no
// data-dependent inner loop (the reference's loop over (a, b] is unbounded for
// sparse values), identical result: offsets[v] = first i with values[i] >= v.
#include "common.cuh"

namespace {

template <typename W, typename I, bool GATHER>
__global__ void __launch_bounds__(256)
index_kernel(const W* __restrict__ in, const I* __restrict__ idx, W* __restrict__ out, uint64_t n_words,
             uint32_t words_per_value) {
    for (uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; w < n_words;
         w += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t i = w / words_per_value, k = w - i * words_per_value;
        const uint64_t j = (uint64_t)idx[i];
        if (GATHER) out[w] = in[j * words_per_value + k];
        else out[j * words_per_value + k] = in[w];
    }
}

template <typename I, bool GATHER>
int launch_index(const void* in, const void* idx, void* out, uint64_t n, uint32_t value_bytes, cudaStream_t s) {
    if (n == 0) return 0;
    const uintptr_t al = reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out);
#define GO(W)                                                                                            \
    {                                                                                                    \
        const uint32_t wpv = value_bytes / sizeof(W);                                                    \
        const uint64_t nw = n * wpv;                                                                     \
        unsigned blocks = col_blocks(nw, 256);                                                           \
        if (blocks > COL_NUM_SMS * 32) blocks = COL_NUM_SMS * 32;                                        \
        index_kernel<W, I, GATHER><<<blocks, 256, 0, s>>>((const W*)in, (const I*)idx, (W*)out, nw, wpv); \
    }
    if (value_bytes % 16 == 0 && al % 16 == 0) GO(uint4)
    else if (value_bytes % 8 == 0 && al % 8 == 0) GO(uint64_t)
    else if (value_bytes % 4 == 0 && al % 4 == 0) GO(uint32_t)
    else if (value_bytes % 2 == 0 && al % 2 == 0) GO(uint16_t)
    else GO(uint8_t)
#undef GO
    COL_CHECK_LAUNCH();
    return 0;
}

template <typename V, typename O>
__global__ void __launch_bounds__(256)
find_offsets_kernel(const V* __restrict__ values, uint64_t n_values, O* __restrict__ offsets, uint64_t n_offsets) {
    for (uint64_t v = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; v < n_offsets;
         v += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t lo = 0, hi = n_values;  // first i with values[i] >= v
        while (lo < hi) {
            const uint64_t mid = (lo + hi) >> 1;
            if ((uint64_t)__ldg(values + mid) < v) lo = mid + 1;
            else hi = mid;
        }
        // The reference zeroes the entries v <= values[0] from work-item v of an
        // (n_values - 1)-wide launch (offset.cl:10-11), so entries with
        // v >= n_values - 1 keep the host fill.  Reproduced for bit parity.
        if (lo == 0 && v + 1 >= n_values) lo = n_values;
        offsets[v] = (O)lo;
    }
}

__global__ void __launch_bounds__(256) fill_u32_kernel(uint32_t* dst, uint32_t value, uint64_t n) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        dst[i] = value;
}

__global__ void __launch_bounds__(256) range_kernel(uint32_t* ids, uint64_t n) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        ids[i] = (uint32_t)i;
}

}  // namespace

COL_API int col_range(uint32_t* ids, uint64_t n, void* stream) {
    if (!ids) return COL_ERR_BAD_ARG;
    if (n == 0) return 0;
    unsigned blocks = col_blocks(n, 256);
    if (blocks > COL_NUM_SMS * 32) blocks = COL_NUM_SMS * 32;
    range_kernel<<<blocks, 256, 0, col_stream(stream)>>>(ids, n);
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_copy(void* dst, const void* src, uint64_t bytes, void* stream) {
    if (!dst || !src) return COL_ERR_BAD_ARG;
    if (bytes == 0) return 0;
    return (int)cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, col_stream(stream));
}

COL_API int col_gather(const void* in, const void* idx, void* out, uint64_t n, uint32_t value_bytes,
                       uint32_t index_bytes, void* stream) {
    if (!in || !idx || !out || value_bytes == 0) return COL_ERR_BAD_ARG;
    if (index_bytes == 4) return launch_index<uint32_t, true>(in, idx, out, n, value_bytes, col_stream(stream));
    if (index_bytes == 8) return launch_index<uint64_t, true>(in, idx, out, n, value_bytes, col_stream(stream));
    return COL_ERR_UNSUPPORTED;
}

COL_API int col_scatter(const void* in, const void* idx, void* out, uint64_t n, uint32_t value_bytes,
                        uint32_t index_bytes, void* stream) {
    if (!in || !idx || !out || value_bytes == 0) return COL_ERR_BAD_ARG;
    if (index_bytes == 4) return launch_index<uint32_t, false>(in, idx, out, n, value_bytes, col_stream(stream));
    if (index_bytes == 8) return launch_index<uint64_t, false>(in, idx, out, n, value_bytes, col_stream(stream));
    return COL_ERR_UNSUPPORTED;
}

COL_API int col_find_offsets(const void* values, uint64_t n_values, uint32_t value_bytes, void* offsets,
                             uint64_t n_offsets, uint32_t offset_bytes, void* stream) {
    if (!values || !offsets) return COL_ERR_BAD_ARG;
    if (n_offsets == 0) return 0;
    unsigned blocks = col_blocks(n_offsets, 256);
    if (blocks > COL_NUM_SMS * 32) blocks = COL_NUM_SMS * 32;
    cudaStream_t s = col_stream(stream);
#define GO(V, O) find_offsets_kernel<V, O><<<blocks, 256, 0, s>>>((const V*)values, n_values, (O*)offsets, n_offsets)
    if (value_bytes == 4 && offset_bytes == 4) GO(uint32_t, uint32_t);
    else if (value_bytes == 4 && offset_bytes == 8) GO(uint32_t, uint64_t);
    else if (value_bytes == 8 && offset_bytes == 4) GO(uint64_t, uint32_t);
    else if (value_bytes == 8 && offset_bytes == 8) GO(uint64_t, uint64_t);
    else return COL_ERR_UNSUPPORTED;
#undef GO
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API int col_fill_u32(uint32_t* dst, uint32_t value, uint64_t n, void* stream) {
    if (!dst) return COL_ERR_BAD_ARG;
    if (n == 0) return 0;
    unsigned blocks = col_blocks(n, 256);
    if (blocks > COL_NUM_SMS * 32) blocks = COL_NUM_SMS * 32;
    fill_u32_kernel<<<blocks, 256, 0, col_stream(stream)>>>(dst, value, n);
    COL_CHECK_LAUNCH();
    return 0;
}
