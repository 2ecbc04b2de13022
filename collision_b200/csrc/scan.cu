// scan.cu -- exclusive prefix sum of uint32: single pass with decoupled look-back, and from 2^23
// values on a persistent chunked grid that re-reads each tile from L2 (second half of the file).
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
#include <cstdio>
#include <mutex>

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
// From COL_SCAN_LAGGED_MIN values on, the look-back latency of the single-pass kernel costs more
// than it saves (measured: half of copy speed).  A persistent grid (every block co-resident:
// cooperative launch) cuts the array into CHUNKS of one tile per block, and every block alternates:
// sum my tile of chunk c and publish the sum (phase 1), then scan my tile of chunk c - LAG in
// place (phase 2).  By then every block's sum of that chunk has long been published -- there is
// no chain of look-backs: a block adds the `grid` sums of the chunk itself, all loads in flight
// at once -- and the tile is still in L2 (LAG + 1 chunks, ~45 MB of the 126, are live), so the
// second read costs no HBM traffic: 8 B per value.  A sum is a 64-bit word {valid:1, sum:63};
// the words are zeroed before the launch.  Waits only ever point at lower chunks, and phase 1
// never waits, so the grid cannot deadlock.  Measured alternatives (profiles/r1g_scan_variants*.txt):
// LAG 1 (blocks wait on stragglers, 1.5x slower), LAG 3-4 and L2 prefetch of later chunks (the
// live set outgrows what L2 keeps), 16 KB tiles x 4 blocks per SM (-3 %), issuing the re-read before
// phase 1 (registers), and plain reduce-then-scan over one segment per block (12 B per value).
#ifndef COL_SCAN_LAGGED_MIN
#define COL_SCAN_LAGGED_MIN (1ull << 23)   // measured crossover (profiles/r1g_scan_threshold.txt)
#endif
#ifndef COL_SCAN_LAGGED
#define COL_SCAN_LAGGED 1
#endif
#ifndef COL_SCAN_LG_VECS
#define COL_SCAN_LG_VECS 8
#endif
#ifndef COL_SCAN_LG_LAG
#define COL_SCAN_LG_LAG 2
#endif
#ifndef COL_SCAN_LG_CTAS
#define COL_SCAN_LG_CTAS 3            // blocks per SM (80 registers per thread)
#endif
constexpr int LG_THREADS = 256;
constexpr int LG_VECS = COL_SCAN_LG_VECS;         // uint4 per thread per tile
constexpr int LG_TILE = LG_THREADS * LG_VECS * 4; // 8192 values
constexpr int LG_WARP_ITEMS = 32 * LG_VECS * 4;   // contiguous values per warp
constexpr int LG_LAG = COL_SCAN_LG_LAG;
constexpr int LG_MIN_TILE = 2048;                 // scratch is sized for tiles at least this long
constexpr int LG_MAX_GRID = 4096;                 // ... and for grids up to this many blocks
constexpr uint64_t LG_VALID = 1ull << 63;

__global__ void __launch_bounds__(LG_THREADS, COL_SCAN_LG_CTAS)
scan_lagged_kernel(uint32_t* __restrict__ data, uint64_t n, uint32_t chunks, uint64_t* total_out, uint64_t* sums) {
    __shared__ uint64_t s_red[2][LG_THREADS / 32];
    __shared__ uint64_t s_before[LG_THREADS / 32], s_all[LG_THREADS / 32];
    __shared__ uint32_t s_wtot[LG_THREADS / 32];
    const unsigned warp = threadIdx.x >> 5, lane = lane_id();
    const unsigned G = gridDim.x, b = blockIdx.x;
    const uint32_t my_off = warp * LG_WARP_ITEMS + lane * 4;
    uint64_t carry = 0;                            // sum of all chunks before the one being scanned

    for (uint32_t c = 0; c < chunks + LG_LAG; c++) {
        // ---- phase 1: sum of my tile of chunk c
        if (c < chunks) {
            const uint64_t base = ((uint64_t)c * G + b) * LG_TILE;
            uint64_t acc = 0;
            if (base + LG_TILE <= n) {
                uint4 a[LG_VECS];
#pragma unroll
                for (int k = 0; k < LG_VECS; k++) a[k] = *reinterpret_cast<const uint4*>(data + base + my_off + k * 128);
#pragma unroll
                for (int k = 0; k < LG_VECS; k++) acc += (uint64_t)a[k].x + a[k].y + a[k].z + a[k].w;
            } else {
                for (uint64_t e = base + threadIdx.x; e < n; e += LG_THREADS) acc += data[e];
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0) s_red[c & 1][warp] = acc;
            __syncthreads();
            if (threadIdx.x == 0) {
                uint64_t t = 0;
#pragma unroll
                for (int w = 0; w < LG_THREADS / 32; w++) t += s_red[c & 1][w];
                st_relaxed_u64(sums + (uint64_t)c * G + b, LG_VALID | t);
            }
        }
        // ---- phase 2: scan of my tile of chunk c - LAG
        if (c >= LG_LAG) {
            const uint32_t cc = c - LG_LAG;
            const uint64_t base = ((uint64_t)cc * G + b) * LG_TILE;
            const bool full = base + LG_TILE <= n;
            uint4 v[LG_VECS];
#pragma unroll
            for (int k = 0; k < LG_VECS; k++) {       // (L2 hits) issued before the sums are waited for
                const uint64_t e = base + my_off + k * 128;
                if (full) {
                    v[k] = *reinterpret_cast<const uint4*>(data + e);
                } else {
                    v[k].x = e + 0 < n ? data[e + 0] : 0u;
                    v[k].y = e + 1 < n ? data[e + 1] : 0u;
                    v[k].z = e + 2 < n ? data[e + 2] : 0u;
                    v[k].w = e + 3 < n ? data[e + 3] : 0u;
                }
            }
            uint64_t before = 0, all = 0;
            for (unsigned i = threadIdx.x; i < G; i += LG_THREADS) {
                const uint64_t* w = sums + (uint64_t)cc * G + i;
                uint64_t t = ld_relaxed_u64(w);
                while (!(t & LG_VALID)) { __nanosleep(32); t = ld_relaxed_u64(w); }
                t &= ~LG_VALID;
                all += t;
                if (i < b) before += t;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                before += __shfl_xor_sync(0xffffffffu, before, o);
                all += __shfl_xor_sync(0xffffffffu, all, o);
            }
            if (lane == 0) { s_before[warp] = before; s_all[warp] = all; }
            // per-row warp scans of the tile meanwhile (mod 2^32)
            uint32_t excl[LG_VECS];
            uint32_t row_base = 0;
#pragma unroll
            for (int k = 0; k < LG_VECS; k++) {
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
            if (lane == 0) s_wtot[warp] = row_base;
            __syncthreads();
            uint64_t bsum = 0, asum = 0;
            uint32_t warp_prefix = 0;
#pragma unroll
            for (int w = 0; w < LG_THREADS / 32; w++) {
                bsum += s_before[w];
                asum += s_all[w];
                if (w < (int)warp) warp_prefix += s_wtot[w];
            }
            const uint32_t b0 = (uint32_t)(carry + bsum) + warp_prefix;
            carry += asum;
#pragma unroll
            for (int k = 0; k < LG_VECS; k++) {
                const uint64_t e = base + my_off + k * 128;
                uint4 o;
                o.x = b0 + excl[k];
                o.y = o.x + v[k].x;
                o.z = o.y + v[k].y;
                o.w = o.z + v[k].z;
                if (full) {
                    *reinterpret_cast<uint4*>(data + e) = o;
                } else {
                    if (e + 0 < n) data[e + 0] = o.x;
                    if (e + 1 < n) data[e + 1] = o.y;
                    if (e + 2 < n) data[e + 2] = o.z;
                    if (e + 3 < n) data[e + 3] = o.w;
                }
            }
            __syncthreads();     // s_before / s_all / s_wtot are rewritten by the next round
        }
    }
    if (b == 0 && threadIdx.x == 0 && total_out) *total_out = carry;
}

int lagged_grid() {
    int dev = 0, sms = 0, per_sm = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, scan_lagged_kernel, LG_THREADS, 0) != cudaSuccess)
        return 0;
    if (per_sm > COL_SCAN_LG_CTAS) per_sm = COL_SCAN_LG_CTAS;
    const int g = sms * per_sm;
    return g > LG_MAX_GRID ? LG_MAX_GRID : g;
}

}  // namespace

COL_API size_t col_scan_scratch_bytes(uint64_t n) {
    // one 64-bit word per tile of whichever kernel runs (the shortest tiles are the chunked kernel's, which
    // also rounds the tile count up to whole chunks of at most LG_MAX_GRID tiles)
    static_assert(LG_TILE >= LG_MIN_TILE && SCAN_TILE >= LG_MIN_TILE, "scratch");
    return SCAN_STATUS_OFF + (size_t)(n / LG_MIN_TILE + 1 + LG_MAX_GRID) * sizeof(uint64_t);
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
    if (COL_SCAN_LAGGED && n >= COL_SCAN_LAGGED_MIN && (reinterpret_cast<uintptr_t>(data) & 15u) == 0) {
        const int grid = lagged_grid();
        if (grid > 0) {
            const uint64_t per_chunk = (uint64_t)grid * LG_TILE;
            uint32_t chunks = (uint32_t)((n + per_chunk - 1) / per_chunk);
            uint64_t* sums = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(scratch) + SCAN_STATUS_OFF);
            COL_CUDA(cudaMemsetAsync(sums, 0, (size_t)chunks * grid * sizeof(uint64_t), s));
            void* args[] = {&data, &n, &chunks, &total_out, &sums};
            // cooperative: the launch fails instead of hanging if the grid could not be co-resident
            COL_CUDA(cudaLaunchCooperativeKernel((const void*)scan_lagged_kernel, dim3((unsigned)grid),
                                                 dim3(LG_THREADS), args, 0, s));
            return 0;
        }
    }
    COL_CUDA(cudaMemsetAsync(scratch, 0, SCAN_STATUS_OFF + (size_t)tiles * sizeof(uint64_t), s));
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
    uint64_t* status = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(scratch) + SCAN_STATUS_OFF);
    scan_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, s>>>(data, n, total_out, ticket, status);
    COL_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------- no-allocation variant
// col_exclusive_scan_u32_static: the single-pass look-back scan above with its ticket and status
// words in a file-scope device array instead of caller scratch, for callers that cannot allocate
// (CUDA graph capture, serving loops).  No counterpart in the reference.
// Every GPU that loads the library holds the array until exit; every call on
// a device shares it, so calls that overlap on that GPU (two streams, two threads) race.
// The array keeps the previous call's flags: the part this call uses is cleared first, in stream order.
namespace {

constexpr uint64_t SCAN_STATIC_TILES = (COL_SCAN_STATIC_MAX_N + SCAN_TILE - 1) / SCAN_TILE;
constexpr size_t SCAN_STATIC_BYTES = SCAN_STATUS_OFF + SCAN_STATIC_TILES * sizeof(uint64_t);   // 16,640 B

static __device__ __align__(256) unsigned char g_scan_static[SCAN_STATIC_BYTES];

constexpr int SCAN_STATIC_MAX_DEVICES = 64;
std::mutex g_scan_static_mutex;
void* g_scan_static_ptr[SCAN_STATIC_MAX_DEVICES];   // device address of g_scan_static, per device

// In host code `g_scan_static` is only a handle for the runtime, not a device address: resolve it
// once per device, with that device current.
int scan_static_address(void** out) {
    int dev = -1;
    COL_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= SCAN_STATIC_MAX_DEVICES) return COL_ERR_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_scan_static_mutex);
    if (!g_scan_static_ptr[dev]) {
        void* p = nullptr;
        COL_CUDA(cudaGetSymbolAddress(&p, g_scan_static));
        if (!p) return (int)cudaErrorInvalidSymbol;
        g_scan_static_ptr[dev] = p;
    }
    *out = g_scan_static_ptr[dev];
    return 0;
}

}  // namespace

COL_API int col_exclusive_scan_u32_static(uint32_t* data, uint64_t n, uint64_t* total_out, void* stream) {
    if (!data && n) return COL_ERR_BAD_ARG;
    if (n > COL_SCAN_STATIC_MAX_N) {
        fprintf(stderr, "col_exclusive_scan_u32_static: n = %llu exceeds COL_SCAN_STATIC_MAX_N = %llu "
                        "(the built-in scratch array); use col_exclusive_scan_u32 with caller scratch\n",
                (unsigned long long)n, (unsigned long long)COL_SCAN_STATIC_MAX_N);
        return COL_ERR_TOO_LARGE;
    }
    cudaStream_t s = col_stream(stream);
    if (n == 0) {
        if (total_out) COL_CUDA(cudaMemsetAsync(total_out, 0, sizeof(uint64_t), s));
        return 0;
    }
    void* scratch = nullptr;
    const int rc = scan_static_address(&scratch);
    if (rc) return rc;
    const uint64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    COL_CUDA(cudaMemsetAsync(scratch, 0, SCAN_STATUS_OFF + (size_t)tiles * sizeof(uint64_t), s));
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
    uint64_t* status = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(scratch) + SCAN_STATUS_OFF);
    scan_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, s>>>(data, n, total_out, ticket, status);
    COL_CHECK_LAUNCH();
    return 0;
}
