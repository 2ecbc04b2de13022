// scan.cu -- single-pass exclusive prefix sum of uint32 (decoupled look-back).
//
// Replaces PrefixScanner.prefix_sum (collision/scan.py:75-112): the reference
// runs 2*levels+1 launches (local_scan per level, then block_scan per level:
// collision/scan.cl:5-36, Blelloch sweeps of collision/local_scan.cl:2-25) and
// touches the data twice.  Here every tile of 8192 values is read once and
// written once (8 B per value, the HBM roofline of the op); the running prefix
// travels between tiles through a 64-bit status word per tile
// {flag:2, aggregate:62}.  Tiles take their index from an atomic ticket so a
// tile can only wait on tiles that already started (forward progress).
// Results wrap mod 2^32 exactly like the reference's uint arithmetic; the u64
// grand total (62-bit) is produced for the pair compaction.
#include "common.cuh"

namespace {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_VECS = 8;                               // uint4 loads per thread (4 -> 8: +23 % at 2^26 values)
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_VECS * 4;    // 8192 values
constexpr int SCAN_WARP_ITEMS = 32 * SCAN_VECS * 4;        // 1024 values per warp
constexpr size_t SCAN_STATUS_OFF = 256;

constexpr uint64_t FLAG_PARTIAL = 1ull << 62;
constexpr uint64_t FLAG_INCLUSIVE = 2ull << 62;
constexpr uint64_t VALUE_MASK = (1ull << 62) - 1;

__global__ void __launch_bounds__(SCAN_THREADS)
scan_kernel(uint32_t* __restrict__ data, uint64_t n, uint64_t* total_out, unsigned* ticket, uint64_t* status) {
    __shared__ unsigned s_tile;
    __shared__ uint64_t s_warp_total[SCAN_THREADS / 32];
    __shared__ uint64_t s_tile_prefix;

    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const unsigned tile = s_tile;
    const uint64_t tile_base = (uint64_t)tile * SCAN_TILE;
    const unsigned warp = threadIdx.x >> 5, lane = lane_id();
    const uint64_t warp_base = tile_base + (uint64_t)warp * SCAN_WARP_ITEMS;
    const bool full = tile_base + SCAN_TILE <= n;
    const bool vec_ok = (reinterpret_cast<uintptr_t>(data) & 15u) == 0;

    uint4 v[SCAN_VECS];
#pragma unroll
    for (int k = 0; k < SCAN_VECS; k++) {
        const uint64_t e = warp_base + (uint64_t)(k * 32 + lane) * 4;
        if (full && vec_ok) {
            v[k] = *reinterpret_cast<const uint4*>(data + e);
        } else {
            v[k].x = e + 0 < n ? data[e + 0] : 0u;
            v[k].y = e + 1 < n ? data[e + 1] : 0u;
            v[k].z = e + 2 < n ? data[e + 2] : 0u;
            v[k].w = e + 3 < n ? data[e + 3] : 0u;
        }
    }

    // per-row (k) inclusive warp scan of the thread sums, in u64 so that the
    // tile aggregate (and the grand total) does not wrap
    uint64_t excl[SCAN_VECS];
    uint64_t row_base = 0;
#pragma unroll
    for (int k = 0; k < SCAN_VECS; k++) {
        const uint64_t s = (uint64_t)v[k].x + v[k].y + v[k].z + v[k].w;
        uint64_t inc = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint64_t up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += up;
        }
        excl[k] = row_base + inc - s;
        row_base += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) s_warp_total[warp] = row_base;
    __syncthreads();

    uint64_t warp_prefix = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; w++) {
        const uint64_t t = s_warp_total[w];
        if (w < warp) warp_prefix += t;
        tile_total += t;
    }

    // decoupled look-back by warp 0 (32 predecessors per step)
    if (warp == 0) {
        uint64_t prefix = 0;
        if (tile == 0) {
            if (lane == 0) st_relaxed_u64(status, FLAG_INCLUSIVE | (tile_total & VALUE_MASK));
        } else {
            if (lane == 0) st_relaxed_u64(status + tile, FLAG_PARTIAL | (tile_total & VALUE_MASK));
            int64_t look = (int64_t)tile - 1 - lane;
            while (true) {
                uint64_t st = FLAG_INCLUSIVE;  // lanes before tile 0 read as an empty inclusive prefix
                if (look >= 0) {
                    do { st = ld_relaxed_u64(status + look); } while ((st >> 62) == 0);
                }
                const unsigned inc_mask = __ballot_sync(0xffffffffu, (st >> 62) == 2);
                // lanes up to and including the nearest inclusive predecessor contribute
                const unsigned first = inc_mask ? (unsigned)__ffs(inc_mask) - 1 : 31u;
                uint64_t c = (lane <= first) ? (st & VALUE_MASK) : 0;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                prefix += c;
                if (inc_mask) break;
                look -= 32;
            }
            if (lane == 0) st_relaxed_u64(status + tile, FLAG_INCLUSIVE | ((prefix + tile_total) & VALUE_MASK));
        }
        if (lane == 0) {
            s_tile_prefix = prefix;
            if (total_out && tile_base + SCAN_TILE >= n) *total_out = prefix + tile_total;
        }
    }
    __syncthreads();
    const uint32_t base = (uint32_t)(s_tile_prefix + warp_prefix);

#pragma unroll
    for (int k = 0; k < SCAN_VECS; k++) {
        const uint64_t e = warp_base + (uint64_t)(k * 32 + lane) * 4;
        uint4 o;
        o.x = base + (uint32_t)excl[k];
        o.y = o.x + v[k].x;
        o.z = o.y + v[k].y;
        o.w = o.z + v[k].z;
        if (full && vec_ok) {
            *reinterpret_cast<uint4*>(data + e) = o;
        } else {
            if (e + 0 < n) data[e + 0] = o.x;
            if (e + 1 < n) data[e + 1] = o.y;
            if (e + 2 < n) data[e + 2] = o.z;
            if (e + 3 < n) data[e + 3] = o.w;
        }
    }
}

}  // namespace

COL_API size_t col_scan_scratch_bytes(uint64_t n) {
    const uint64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    return SCAN_STATUS_OFF + (size_t)(tiles ? tiles : 1) * sizeof(uint64_t);
}

COL_API int col_exclusive_scan_u32(uint32_t* data, uint64_t n, uint64_t* total_out, void* scratch, void* stream) {
    if (!scratch || (!data && n)) return COL_ERR_BAD_ARG;
    cudaStream_t s = col_stream(stream);
    if (n == 0) {
        if (total_out) COL_CUDA(cudaMemsetAsync(total_out, 0, sizeof(uint64_t), s));
        return 0;
    }
    const uint64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (tiles > 0x7fffffffull) return COL_ERR_TOO_LARGE;
    COL_CUDA(cudaMemsetAsync(scratch, 0, col_scan_scratch_bytes(n), s));
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
    uint64_t* status = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(scratch) + SCAN_STATUS_OFF);
    scan_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, s>>>(data, n, total_out, ticket, status);
    COL_CHECK_LAUNCH();
    return 0;
}
