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

// ---------------------------------------------------------------- large arrays
// Above COL_SCAN_TWO_PHASE_MIN values the look-back latency of the single-pass kernel costs more
// than a second read: a persistent grid (every block co-resident, cooperative launch) cuts the
// array into one contiguous segment per block and runs reduce-then-scan with ONE grid-wide
// arrival barrier in between.  Phase 1 sums the segment BACK TO FRONT, phase 2 scans it front to
// back, in place: the part of the segment read last is re-read first, while it still sits in
// the 126 MB L2, and the scan's stores land on lines it has just read.  HBM traffic is
// 12 B per value minus the L2 hits of the re-read (8 B algorithmic); no block ever waits on
// another except at the barrier.
#ifndef COL_SCAN_TWO_PHASE_MIN
#define COL_SCAN_TWO_PHASE_MIN (1ull << 24)
#endif
#ifndef COL_SCAN_RS_VECS
#define COL_SCAN_RS_VECS 4
#endif
#ifndef COL_SCAN_RS_MAX_CTAS
#define COL_SCAN_RS_MAX_CTAS 4
#endif
constexpr int RS_THREADS = 256;
constexpr int RS_VECS = COL_SCAN_RS_VECS;                 // uint4 per thread per tile
constexpr int RS_TILE = RS_THREADS * RS_VECS * 4;         // values per tile
constexpr int RS_WARP_ITEMS = 32 * RS_VECS * 4;           // contiguous values per warp per tile
constexpr size_t RS_TOTALS_OFF = 256;                     // scratch: [0] arrival counter, then u64 per block

__global__ void __launch_bounds__(RS_THREADS)
scan_two_phase_kernel(uint32_t* __restrict__ data, uint64_t n, uint64_t seg_len, uint64_t* total_out,
                      uint32_t* arrivals, uint64_t* seg_total) {
    __shared__ uint64_t s_red[RS_THREADS / 32];
    __shared__ uint64_t s_pre[2];
    __shared__ uint32_t s_wtot[2][RS_THREADS / 32];
    const unsigned warp = threadIdx.x >> 5, lane = lane_id();
    const uint64_t seg_begin = min(n, (uint64_t)blockIdx.x * seg_len);
    const uint64_t seg_end = min(n, seg_begin + seg_len);
    const int64_t tiles = (int64_t)((seg_end - seg_begin + RS_TILE - 1) / RS_TILE);
    // element offset, inside a tile, of this thread's k-th vector (warp-major, as in scan_kernel)
    const uint32_t my_off = warp * RS_WARP_ITEMS + lane * 4;

    // ---- phase 1: segment total, back to front
    uint64_t acc = 0;
    for (int64_t t = tiles - 1; t >= 0; t--) {
        const uint64_t base = seg_begin + (uint64_t)t * RS_TILE;
        if (base + RS_TILE <= seg_end) {
            uint4 v[RS_VECS];
#pragma unroll
            for (int k = 0; k < RS_VECS; k++) v[k] = *reinterpret_cast<const uint4*>(data + base + my_off + k * 128);
#pragma unroll
            for (int k = 0; k < RS_VECS; k++) acc += (uint64_t)v[k].x + v[k].y + v[k].z + v[k].w;
        } else {
            for (uint64_t e = base + threadIdx.x; e < seg_end; e += RS_THREADS) acc += data[e];
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) s_red[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t t = 0;
#pragma unroll
        for (int w = 0; w < RS_THREADS / 32; w++) t += s_red[w];
        st_relaxed_u64(seg_total + blockIdx.x, t);
        __threadfence();
        atomicAdd(arrivals, 1u);
        while (ld_acquire_u32(arrivals) < gridDim.x) __nanosleep(64);
    }
    __syncthreads();

    // ---- prefix of the segment totals before mine (and the grand total)
    {
        uint64_t before = 0, all = 0;
        for (unsigned b = threadIdx.x; b < gridDim.x; b += RS_THREADS) {
            const uint64_t t = ld_relaxed_u64(seg_total + b);
            all += t;
            if (b < blockIdx.x) before += t;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            before += __shfl_xor_sync(0xffffffffu, before, o);
            all += __shfl_xor_sync(0xffffffffu, all, o);
        }
        __syncthreads();     // s_red is reused
        if (lane == 0) s_red[warp] = before;
        if (blockIdx.x == 0 && total_out) {      // block 0 alone needs the grand total
            __shared__ uint64_t s_all[RS_THREADS / 32];
            if (lane == 0) s_all[warp] = all;
            __syncthreads();
            if (threadIdx.x == 0) {
                uint64_t t = 0;
#pragma unroll
                for (int w = 0; w < RS_THREADS / 32; w++) t += s_all[w];
                *total_out = t;
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint64_t t = 0;
#pragma unroll
            for (int w = 0; w < RS_THREADS / 32; w++) t += s_red[w];
            s_pre[0] = t;
        }
        __syncthreads();
    }

    // ---- phase 2: exclusive scan of the segment, front to back, in place (mod 2^32)
    uint32_t carry = (uint32_t)s_pre[0];
    for (int64_t t = 0; t < tiles; t++) {
        const uint64_t base = seg_begin + (uint64_t)t * RS_TILE;
        const bool full = base + RS_TILE <= seg_end;
        uint4 v[RS_VECS];
#pragma unroll
        for (int k = 0; k < RS_VECS; k++) {
            const uint64_t e = base + my_off + k * 128;
            if (full) {
                v[k] = *reinterpret_cast<const uint4*>(data + e);
            } else {
                v[k].x = e + 0 < seg_end ? data[e + 0] : 0u;
                v[k].y = e + 1 < seg_end ? data[e + 1] : 0u;
                v[k].z = e + 2 < seg_end ? data[e + 2] : 0u;
                v[k].w = e + 3 < seg_end ? data[e + 3] : 0u;
            }
        }
        uint32_t excl[RS_VECS];
        uint32_t row_base = 0;
#pragma unroll
        for (int k = 0; k < RS_VECS; k++) {
            const uint32_t sum = v[k].x + v[k].y + v[k].z + v[k].w;
            uint32_t inc = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += up;
            }
            excl[k] = row_base + inc - sum;
            row_base += __shfl_sync(0xffffffffu, inc, 31);
        }
        uint32_t* wt = s_wtot[t & 1];            // double-buffered: one barrier per tile
        if (lane == 0) wt[warp] = row_base;
        __syncthreads();
        uint32_t warp_prefix = 0, tile_total = 0;
#pragma unroll
        for (int w = 0; w < RS_THREADS / 32; w++) {
            const uint32_t x = wt[w];
            if (w < (int)warp) warp_prefix += x;
            tile_total += x;
        }
        const uint32_t b0 = carry + warp_prefix;
#pragma unroll
        for (int k = 0; k < RS_VECS; k++) {
            const uint64_t e = base + my_off + k * 128;
            uint4 o;
            o.x = b0 + excl[k];
            o.y = o.x + v[k].x;
            o.z = o.y + v[k].y;
            o.w = o.z + v[k].z;
            if (full) {
                *reinterpret_cast<uint4*>(data + e) = o;
            } else {
                if (e + 0 < seg_end) data[e + 0] = o.x;
                if (e + 1 < seg_end) data[e + 1] = o.y;
                if (e + 2 < seg_end) data[e + 2] = o.z;
                if (e + 3 < seg_end) data[e + 3] = o.w;
            }
        }
        carry += tile_total;
    }
}

// blocks of the persistent grid: what the device can hold at once (0 = query failed)
int two_phase_grid() {
    int dev = 0, sms = 0, per_sm = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, scan_two_phase_kernel, RS_THREADS, 0) != cudaSuccess)
        return 0;
    if (per_sm > COL_SCAN_RS_MAX_CTAS) per_sm = COL_SCAN_RS_MAX_CTAS;
    return sms * per_sm;
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
    if (n >= COL_SCAN_TWO_PHASE_MIN && (reinterpret_cast<uintptr_t>(data) & 15u) == 0) {
        const int grid = two_phase_grid();
        // the per-block totals take the place of the tile status words (far fewer blocks than tiles here)
        if (grid > 0 && RS_TOTALS_OFF + (size_t)grid * sizeof(uint64_t) <= col_scan_scratch_bytes(n)) {
            COL_CUDA(cudaMemsetAsync(scratch, 0, sizeof(uint32_t), s));
            uint32_t* arrivals = reinterpret_cast<uint32_t*>(scratch);
            uint64_t* seg_total = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(scratch) + RS_TOTALS_OFF);
            uint64_t seg_len = (n + grid - 1) / grid;
            seg_len = (seg_len + RS_TILE - 1) / RS_TILE * RS_TILE;     // whole tiles: segments stay 16-byte aligned
            void* args[] = {&data, &n, &seg_len, &total_out, &arrivals, &seg_total};
            // cooperative: the launch fails instead of hanging if the grid could not be co-resident
            COL_CUDA(cudaLaunchCooperativeKernel((const void*)scan_two_phase_kernel, dim3((unsigned)grid),
                                                 dim3(RS_THREADS), args, 0, s));
            return 0;
        }
    }
    COL_CUDA(cudaMemsetAsync(scratch, 0, col_scan_scratch_bytes(n), s));
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
    uint64_t* status = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(scratch) + SCAN_STATUS_OFF);
    scan_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, s>>>(data, n, total_out, ticket, status);
    COL_CHECK_LAUNCH();
    return 0;
}
