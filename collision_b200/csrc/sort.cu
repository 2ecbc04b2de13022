// sort.cu -- stable LSD radix sort, one-sweep with decoupled look-back.
//
// Replaces RadixSorter.sort (collision/radix.py:118-170).  The reference runs,
// for each 4-bit digit, block_sort (in-place local-memory sort by successive
// 1-bit splits + digit-major histogram, collision/radix.cl:48-102), a
// multi-launch prefix scan of that histogram, scatter (radix.cl:104-139) and two
// full-array copies back -- ~384 B moved per (key, value).  Here:
//
//   1. ONE histogram kernel reads the keys once and counts every 8-bit digit of
//      every pass into shared-memory histograms (4 x 256 bins for u32 keys).
//   2. A one-block kernel turns the histograms into exclusive digit bases.
//   3. ONE kernel per pass: a tile of keys is loaded warp-striped, ranked inside
//      its warp with match.any (stable: lane order within a row, rows in
//      order), combined across warps through per-warp digit counters in shared
//      memory, the tile's 256 digit counts are chained to the preceding tiles by
//      decoupled look-back on 32-bit {flag:2, count:30} words, and the tile is
//      written out through shared memory in digit order so that global stores
//      are runs of consecutive addresses.
//
// Traffic: 4n + passes * 2 * (key + value) * n bytes -> 68 B per (u32, u32)
// pair, 36 B per u32 key: the HBM roofline of the op.  Stability is part of
// the contract (the sorted position is the leaf index of the tree).
#include <cstdlib>

#include "common.cuh"
#include "sort_col.cuh"

namespace {

constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int MAX_PASSES = 8;

#ifndef COL_SORT_THREADS
#define COL_SORT_THREADS 512
#endif
#ifndef COL_SORT_CTAS
#define COL_SORT_CTAS 2
#endif
constexpr int SORT_THREADS = COL_SORT_THREADS;
constexpr int SORT_WARPS = SORT_THREADS / 32;

constexpr uint32_t ST_PARTIAL = 1u << 30;
constexpr uint32_t ST_INCLUSIVE = 2u << 30;
constexpr uint32_t ST_VALUE = (1u << 30) - 1;

template <typename K> struct Tile { static constexpr int ITEMS = 16; };
template <> struct Tile<uint64_t> { static constexpr int ITEMS = 8; };

// ---------------------------------------------------------------- histogram
constexpr int HIST_THREADS = 512;
constexpr int HIST_ITEMS = 16;  // keys per thread per iteration

template <typename K>
__global__ void __launch_bounds__(HIST_THREADS)
histogram_kernel(const K* __restrict__ keys, uint64_t n, uint32_t* __restrict__ ghist, uint32_t shift0,
                 uint32_t bits, uint32_t npass) {
    __shared__ uint32_t sh[MAX_PASSES * RADIX];
    for (int i = threadIdx.x; i < MAX_PASSES * RADIX; i += HIST_THREADS) sh[i] = 0;
    __syncthreads();
    const uint32_t mask = (1u << bits) - 1;
    const uint64_t chunk = (uint64_t)HIST_THREADS * HIST_ITEMS;
    for (uint64_t base = (uint64_t)blockIdx.x * chunk; base < n; base += (uint64_t)gridDim.x * chunk) {
        K k[HIST_ITEMS];
#pragma unroll
        for (int i = 0; i < HIST_ITEMS; i++) {
            const uint64_t e = base + (uint64_t)i * HIST_THREADS + threadIdx.x;
            k[i] = e < n ? keys[e] : (K)0;
        }
#pragma unroll
        for (int i = 0; i < HIST_ITEMS; i++) {
            const uint64_t e = base + (uint64_t)i * HIST_THREADS + threadIdx.x;
            if (e < n)
                for (uint32_t p = 0; p < npass; p++)
                    atomicAdd(&sh[p * RADIX + (uint32_t)((k[i] >> (shift0 + p * bits)) & mask)], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < (int)npass * RADIX; i += HIST_THREADS)
        if (sh[i]) atomicAdd(&ghist[i], sh[i]);
}

// exclusive scan of each pass's 256 bins, in place (one warp-shuffle block scan per pass)
__global__ void __launch_bounds__(RADIX) digit_base_kernel(uint32_t* ghist) {
    __shared__ uint32_t warp_tot[RADIX / 32];
    uint32_t* h = ghist + blockIdx.x * RADIX;
    const uint32_t c = h[threadIdx.x];
    uint32_t inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane_id() >= o) inc += up;
    }
    if (lane_id() == 31) warp_tot[threadIdx.x >> 5] = inc;
    __syncthreads();
    uint32_t base = 0;
    for (int w = 0; w < (int)(threadIdx.x >> 5); w++) base += warp_tot[w];
    h[threadIdx.x] = base + inc - c;
}

// ----------------------------------------------------------------- onesweep
// index of element e of the sorted tile in shared memory: one pad element after every 32
__host__ __device__ constexpr uint32_t skewed(uint32_t e) { return e + (e >> 5); }

// Instruction budget notes (B200: the integer ALU pipe issues one warp instruction every
// 2 cycles per scheduler, POPC/FLO one every 8 -- the ranking below is written against
// that, not against the issue rate):
//  * peers of a key = lanes of its row with the same digit: one VOTE per digit bit;
//  * a lane's rank among its peers needs ONE popc (peers below me); the group's size is
//    that popc + 1 on the highest peer, which alone bumps the warp's digit counter;
//  * the counters hold BYTE offsets, and every per-(warp, digit) table sits at a constant
//    displacement from the counter, so one address per key serves every lookup.
// lanes of the warp whose 8-bit digit equals mine: one ballot per digit bit, four
// instructions each (test, vote, select, combine) -- spelled out in PTX because the compiler's
// own rendering of the obvious C loop spends six.
__device__ __forceinline__ unsigned match_digit(uint32_t d) {
    unsigned p;
    asm(
        "{\n"
        ".reg .pred q;\n"
        ".reg .b32 v, m, t;\n"
        "and.b32 t, %1, 1;   setp.ne.u32 q, t, 0; vote.sync.ballot.b32 v, q, 0xffffffff; selp.b32 m, 0, 0xffffffff, q; xor.b32 %0, v, m;\n"
        "and.b32 t, %1, 2;   setp.ne.u32 q, t, 0; vote.sync.ballot.b32 v, q, 0xffffffff; selp.b32 m, 0, 0xffffffff, q; lop3.b32 %0, %0, v, m, 0x60;\n"
        "and.b32 t, %1, 4;   setp.ne.u32 q, t, 0; vote.sync.ballot.b32 v, q, 0xffffffff; selp.b32 m, 0, 0xffffffff, q; lop3.b32 %0, %0, v, m, 0x60;\n"
        "and.b32 t, %1, 8;   setp.ne.u32 q, t, 0; vote.sync.ballot.b32 v, q, 0xffffffff; selp.b32 m, 0, 0xffffffff, q; lop3.b32 %0, %0, v, m, 0x60;\n"
        "and.b32 t, %1, 16;  setp.ne.u32 q, t, 0; vote.sync.ballot.b32 v, q, 0xffffffff; selp.b32 m, 0, 0xffffffff, q; lop3.b32 %0, %0, v, m, 0x60;\n"
        "and.b32 t, %1, 32;  setp.ne.u32 q, t, 0; vote.sync.ballot.b32 v, q, 0xffffffff; selp.b32 m, 0, 0xffffffff, q; lop3.b32 %0, %0, v, m, 0x60;\n"
        "and.b32 t, %1, 64;  setp.ne.u32 q, t, 0; vote.sync.ballot.b32 v, q, 0xffffffff; selp.b32 m, 0, 0xffffffff, q; lop3.b32 %0, %0, v, m, 0x60;\n"
        "and.b32 t, %1, 128; setp.ne.u32 q, t, 0; vote.sync.ballot.b32 v, q, 0xffffffff; selp.b32 m, 0, 0xffffffff, q; lop3.b32 %0, %0, v, m, 0x60;\n"
        "}\n"
        : "=r"(p)
        : "r"(d));
    return p;
}

template <typename K, int ITEMS, bool FULL, bool HAS_VALUES>
__device__ __forceinline__ void onesweep_tile(const K* __restrict__ keys_in, K* __restrict__ keys_out,
                                              const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out,
                                              uint32_t n, const uint32_t* __restrict__ digit_base,
                                              uint32_t* __restrict__ status, uint32_t shift, uint32_t mask,
                                              uint32_t bits, unsigned tile, unsigned char* smem_raw,
                                              uint32_t* s_digit_start, uint32_t* s_global_off, uint32_t* s_scan_warp) {
    constexpr int TILE = SORT_THREADS * ITEMS;
    constexpr int WARP_ITEMS = 32 * ITEMS;
    constexpr uint32_t KB = sizeof(K);
    // the sorted tile lives in shared memory with one pad element after every 32 (skewed()):
    // when the 32 keys of a row all carry different digits with equal counts so far (sorted or
    // strided input: every digit's run then starts on bank 0) the scatter would otherwise
    // serialise 32 ways -- measured 1.8x on a pass over `arange` keys
    K* s_keys = reinterpret_cast<K*>(smem_raw);                                     // skewed(TILE) keys
    uint32_t* s_whist = reinterpret_cast<uint32_t*>(smem_raw + sizeof(K) * skewed(TILE));   // SORT_WARPS x RADIX
    uint32_t* s_vals = s_whist + SORT_WARPS * RADIX;                                // skewed(TILE) values (if any)

    const uint32_t tile_base = tile * (uint32_t)TILE;
    const uint32_t valid = FULL ? (uint32_t)TILE : n - tile_base;
    const unsigned warp = threadIdx.x >> 5, lane = lane_id();
    const uint32_t warp_off = warp * WARP_ITEMS;

    // ---- load (warp-striped: row i of a warp is 32 consecutive keys)
    K key[ITEMS];
    {
        const K* src = keys_in + tile_base + warp_off + lane;
#pragma unroll
        for (int i = 0; i < ITEMS; i++) {
            if (FULL) key[i] = src[i * 32];
            else key[i] = (warp_off + i * 32 + lane) < valid ? src[i * 32] : (K)~(K)0;  // padding sorts last
        }
    }

    // ---- early counts: per-warp digit histograms (byte units) by non-returning shared atomics
    unsigned char* my_hist = reinterpret_cast<unsigned char*>(s_whist + warp * RADIX);
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const uint32_t d = (uint32_t)(key[i] >> shift) & mask;
        atomicAdd(reinterpret_cast<uint32_t*>(my_hist + d * 4), KB);
    }
    // payload loads are issued here so that their latency hides behind the combine phase
    uint32_t val[HAS_VALUES ? ITEMS : 1];
    if (HAS_VALUES) {
        const uint32_t* src = vals_in + tile_base + warp_off + lane;
#pragma unroll
        for (int i = 0; i < ITEMS; i++) {
            if (FULL) val[i] = src[i * 32];
            else val[i] = (warp_off + i * 32 + lane) < valid ? src[i * 32] : 0u;
        }
    }
    __syncthreads();

    // ---- combine warps: tile count per digit -> published at once (later tiles wait on it);
    // exclusive scan over digits and over warps -> s_whist[w][d] = byte offset, in the sorted
    // tile, of the first key of warp w with digit d
    uint32_t real = 0;  // number of real (non-padding) keys of digit threadIdx.x in this tile
    if (threadIdx.x < RADIX) {
        uint32_t padded = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; w++) padded += s_whist[w * RADIX + threadIdx.x];
        // the padding keys of a partial tile all carry the all-ones digit
        real = padded / KB - ((!FULL && threadIdx.x == mask) ? (uint32_t)TILE - valid : 0u);
        st_relaxed_u32(status + (size_t)tile * RADIX + threadIdx.x, (tile == 0 ? ST_INCLUSIVE : ST_PARTIAL) | real);
        uint32_t inc = padded;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += up;
        }
        if (lane == 31) s_scan_warp[warp] = inc;
        s_digit_start[threadIdx.x] = inc - padded;  // exclusive within this warp of digits
    }
    __syncthreads();
    if (threadIdx.x < RADIX) {
        uint32_t base = 0;
        for (int w = 0; w < (int)warp; w++) base += s_scan_warp[w];
        base += s_digit_start[threadIdx.x];
        s_digit_start[threadIdx.x] = base;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; w++) {
            const uint32_t c = s_whist[w * RADIX + threadIdx.x];
            s_whist[w * RADIX + threadIdx.x] = base;
            base += c;
        }
    }
    __syncthreads();

    // ---- rank inside the warp (stable: lane order within a row, rows in order) and move the
    // key (and payload) to its place in the sorted tile.  The counter of (warp, digit) is the
    // byte offset of the next key of that warp and digit: all peers read it, then the highest
    // peer alone advances it.
    const unsigned lt = lanemask_lt();
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const uint32_t d = (uint32_t)(key[i] >> shift) & mask;
        const unsigned peers = match_digit(d);   // (narrower digits just vote on zero bits)
        const uint32_t below = (uint32_t)__popc(peers & lt) * KB;  // bytes of my peers below me
        uint32_t* ctr = reinterpret_cast<uint32_t*>(my_hist + d * 4);
        const uint32_t r = *reinterpret_cast<volatile uint32_t*>(ctr) + below;
        __syncwarp();
        if ((peers >> lane) == 1u) *reinterpret_cast<volatile uint32_t*>(ctr) = r + KB;
        __syncwarp();
        const uint32_t rs = r + ((r / (32 * KB)) * KB);   // byte offset of skewed(r / KB)
        *reinterpret_cast<K*>(reinterpret_cast<unsigned char*>(s_keys) + rs) = key[i];
        if (HAS_VALUES) *reinterpret_cast<uint32_t*>(reinterpret_cast<unsigned char*>(s_vals) + rs / (KB / 4)) = val[i];
    }

    // ---- decoupled look-back.  The tile's counts went out before the ranking, so by now the
    // predecessors' words are there: no spinning to speak of, only latency -- and that is paid
    // once per LB_WIDE predecessors (independent loads in flight), not once per predecessor.
    // Every warp takes part (digit = thread & 255, the two halves of the block redundantly), so
    // no warp reaches the barrier below much later than the others.
    {
        constexpr int LB_WIDE = 4;
        const unsigned dgt = threadIdx.x & (RADIX - 1);
        uint32_t excl = 0;
        int64_t t = (int64_t)tile - 1;
        bool done = tile == 0;
        while (!done) {
            uint32_t st[LB_WIDE];
#pragma unroll
            for (int k = 0; k < LB_WIDE; k++)
                st[k] = t - k >= 0 ? ld_relaxed_u32(status + (size_t)(t - k) * RADIX + dgt) : ST_INCLUSIVE;
#pragma unroll
            for (int k = 0; k < LB_WIDE; k++) {
                if (!done) {
                    while ((st[k] >> 30) == 0) st[k] = ld_relaxed_u32(status + (size_t)(t - k) * RADIX + dgt);
                    excl += st[k] & ST_VALUE;
                    done = (st[k] >> 30) == 2;
                }
            }
            t -= LB_WIDE;
        }
        if (threadIdx.x < RADIX) {
            if (tile > 0) st_relaxed_u32(status + (size_t)tile * RADIX + dgt, ST_INCLUSIVE | (excl + real));
            s_global_off[dgt] = digit_base[dgt] + excl - s_digit_start[dgt] / KB;
        }
    }
    __syncthreads();

    // ---- write out: consecutive threads -> consecutive tile ranks -> runs of consecutive addresses
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        const uint32_t pos = j * SORT_THREADS + threadIdx.x;
        if (FULL || pos < valid) {
            const K k = s_keys[skewed(pos)];
            const uint32_t d = (uint32_t)(k >> shift) & mask;
            const uint32_t dst = s_global_off[d] + pos;
#ifdef COL_DEBUG   // a wrong look-back sum or digit base shows up as a destination outside the array
            if (dst >= n) __trap();
#endif
            keys_out[dst] = k;
            if (HAS_VALUES) vals_out[dst] = s_vals[skewed(pos)];
        }
    }
}

template <typename K, bool HAS_VALUES>
__global__ void __launch_bounds__(SORT_THREADS, COL_SORT_CTAS)
onesweep_kernel(const K* __restrict__ keys_in, K* __restrict__ keys_out, const uint32_t* __restrict__ vals_in,
                uint32_t* __restrict__ vals_out, uint32_t n, const uint32_t* __restrict__ digit_base,
                uint32_t* __restrict__ status, unsigned* ticket, uint32_t shift, uint32_t mask, uint32_t bits) {
    constexpr int ITEMS = Tile<K>::ITEMS;
    constexpr int TILE = SORT_THREADS * ITEMS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_digit_start[RADIX];
    __shared__ uint32_t s_global_off[RADIX];
    __shared__ uint32_t s_scan_warp[RADIX / 32];
    __shared__ unsigned s_tile;

    uint32_t* s_whist = reinterpret_cast<uint32_t*>(smem_raw + sizeof(K) * skewed(TILE));
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = threadIdx.x; i < SORT_WARPS * RADIX; i += SORT_THREADS) s_whist[i] = 0;
    __syncthreads();
    const unsigned tile = s_tile;
    if ((uint64_t)(tile + 1) * TILE <= n)
        onesweep_tile<K, ITEMS, true, HAS_VALUES>(keys_in, keys_out, vals_in, vals_out, n, digit_base, status, shift,
                                                  mask, bits, tile, smem_raw, s_digit_start, s_global_off, s_scan_warp);
    else
        onesweep_tile<K, ITEMS, false, HAS_VALUES>(keys_in, keys_out, vals_in, vals_out, n, digit_base, status, shift,
                                                   mask, bits, tile, smem_raw, s_digit_start, s_global_off, s_scan_warp);
}

template <typename K, bool HAS_VALUES> constexpr size_t onesweep_smem() {
    return sizeof(K) * skewed(SORT_THREADS * Tile<K>::ITEMS) + sizeof(uint32_t) * SORT_WARPS * RADIX +
           (HAS_VALUES ? sizeof(uint32_t) * skewed(SORT_THREADS * Tile<K>::ITEMS) : 0);
}

// ------------------------------------------------------------ scratch layout
struct SortLayout {
    size_t ghist, tickets, status, status_bytes, tmp_keys, tmp_vals, idx_a, idx_b, total;
    uint64_t tiles;
};

inline size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

SortLayout sort_layout(uint64_t n, uint32_t key_bytes, uint32_t value_bytes) {
    SortLayout L;
    const uint64_t tile = (uint64_t)SORT_THREADS * (key_bytes == 8 ? Tile<uint64_t>::ITEMS : Tile<uint32_t>::ITEMS);
    L.tiles = (n + tile - 1) / tile;
    if (L.tiles == 0) L.tiles = 1;
    size_t off = 0;
    L.ghist = off; off += align_up(sizeof(uint32_t) * MAX_PASSES * RADIX);
    L.tickets = off; off += align_up(sizeof(uint32_t) * MAX_PASSES);
    L.status = off;
    L.status_bytes = align_up(sizeof(uint32_t) * RADIX * L.tiles);
    off += L.status_bytes * key_bytes;  // one status array per 8-bit pass
    L.tmp_keys = off; off += align_up((size_t)n * key_bytes);
    const bool direct = value_bytes == 4, indirect = value_bytes != 0 && value_bytes != 4;
    L.tmp_vals = off; if (direct || indirect) off += align_up((size_t)n * 4);
    L.idx_a = off; if (indirect) off += align_up((size_t)n * 4);
    L.idx_b = off; if (indirect) off += align_up((size_t)n * 4);
    L.total = off;
    return L;
}

__global__ void __launch_bounds__(256) iota_kernel(uint32_t* p, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = i;
}

template <typename K>
int run_passes(const K* keys_in, K* keys_out, const uint32_t* vals_in, uint32_t* vals_out, uint32_t n,
               uint32_t shift0, uint32_t bits, uint32_t npass, char* scratch, const SortLayout& L, cudaStream_t s) {
    uint32_t* ghist = reinterpret_cast<uint32_t*>(scratch + L.ghist);
    unsigned* tickets = reinterpret_cast<unsigned*>(scratch + L.tickets);
    K* tmp_keys = reinterpret_cast<K*>(scratch + L.tmp_keys);
    uint32_t* tmp_vals = reinterpret_cast<uint32_t*>(scratch + L.tmp_vals);
    const bool has_vals = vals_in != nullptr;

    // zero histograms, tickets and all status words in one go (contiguous)
    COL_CUDA(cudaMemsetAsync(scratch, 0, L.status + L.status_bytes * npass, s));
    {
        uint64_t want = ((uint64_t)n + HIST_THREADS * HIST_ITEMS - 1) / (HIST_THREADS * HIST_ITEMS);
        unsigned blocks = (unsigned)(want > COL_NUM_SMS * 4 ? COL_NUM_SMS * 4 : want);
        histogram_kernel<K><<<blocks, HIST_THREADS, 0, s>>>(keys_in, n, ghist, shift0, bits, npass);
        COL_CHECK_LAUNCH();
        digit_base_kernel<<<npass, RADIX, 0, s>>>(ghist);
        COL_CHECK_LAUNCH();
    }
    const uint32_t mask = (1u << bits) - 1;
    // COL_SORT_COLUMN=1: 32-bit keys take the column-counter pass (sort_col.cuh: ranking by counting,
    // no votes) when every buffer a pass reads is 32-byte aligned.  Measured on B200 at 2^26 pairs it
    // EQUALS the ballot pass (1.58 ms; both are bound by the L1/shared-memory pipe at ~25 wavefronts per
    // row of 32 keys, ncu: profiles/r2b_ncu_sortcol_summary.txt) and is slower on sorted keys and at
    // 2^20, so the ballot pass stays the default; the alternative is kept for A/B runs.
    auto aligned32 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 31u) == 0; };
    static const bool column = getenv("COL_SORT_COLUMN") != nullptr;
    const bool col = sizeof(K) == 4 && column && aligned32(keys_in) && aligned32(keys_out) && aligned32(tmp_keys) &&
                     (!has_vals || (aligned32(vals_in) && aligned32(vals_out) && aligned32(tmp_vals)));
    for (uint32_t p = 0; p < npass; p++) {
        // the last pass lands in *_out; earlier ones alternate tmp/out backwards from there
        const bool to_out = ((npass - 1 - p) & 1u) == 0;
        const K* kin = p == 0 ? keys_in : (to_out ? tmp_keys : keys_out);
        K* kout = to_out ? keys_out : tmp_keys;
        const uint32_t* vin = p == 0 ? vals_in : (to_out ? tmp_vals : vals_out);
        uint32_t* vout = to_out ? vals_out : tmp_vals;
        uint32_t* status = reinterpret_cast<uint32_t*>(scratch + L.status + L.status_bytes * p);
        if (col) {
            const int rc = has_vals
                ? colsort::launch<true>((const uint32_t*)kin, (uint32_t*)kout, vin, vout, n, ghist + p * RADIX, status,
                                        tickets + p, shift0 + p * bits, mask, (unsigned)L.tiles, s)
                : colsort::launch<false>((const uint32_t*)kin, (uint32_t*)kout, nullptr, nullptr, n, ghist + p * RADIX,
                                         status, tickets + p, shift0 + p * bits, mask, (unsigned)L.tiles, s);
            if (rc) return rc;
            continue;
        }
        if (has_vals) {
            constexpr size_t smem = onesweep_smem<K, true>();
            COL_CUDA(cudaFuncSetAttribute(onesweep_kernel<K, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            onesweep_kernel<K, true><<<(unsigned)L.tiles, SORT_THREADS, smem, s>>>(
                kin, kout, vin, vout, n, ghist + p * RADIX, status, tickets + p, shift0 + p * bits, mask, bits);
        } else {
            constexpr size_t smem = onesweep_smem<K, false>();
            COL_CUDA(cudaFuncSetAttribute(onesweep_kernel<K, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            onesweep_kernel<K, false><<<(unsigned)L.tiles, SORT_THREADS, smem, s>>>(
                kin, kout, nullptr, nullptr, n, ghist + p * RADIX, status, tickets + p, shift0 + p * bits, mask, bits);
        }
        COL_CHECK_LAUNCH();
    }
    return 0;
}

int sort_impl(const void* keys_in, void* keys_out, const void* vals_in, void* vals_out, uint64_t n,
              uint32_t key_bytes, uint32_t value_bytes, uint32_t shift0, uint32_t bits, uint32_t npass,
              void* scratch_v, cudaStream_t s) {
    if (n == 0) return 0;
    if (n > (1ull << 30)) return COL_ERR_TOO_LARGE;
    char* scratch = reinterpret_cast<char*>(scratch_v);
    const SortLayout L = sort_layout(n, key_bytes, value_bytes);
    const bool have_vals = value_bytes != 0 && vals_in && vals_out;
    const bool indirect = have_vals && value_bytes != 4;
    const uint32_t* vin = have_vals ? (const uint32_t*)vals_in : nullptr;
    uint32_t* vout = have_vals ? (uint32_t*)vals_out : nullptr;
    if (indirect) {  // wide payloads: sort (key, index), then gather the payload once
        uint32_t* ia = reinterpret_cast<uint32_t*>(scratch + L.idx_a);
        iota_kernel<<<col_blocks(n, 256) > COL_NUM_SMS * 16 ? COL_NUM_SMS * 16 : col_blocks(n, 256), 256, 0, s>>>(ia, (uint32_t)n);
        COL_CHECK_LAUNCH();
        vin = ia;
        vout = reinterpret_cast<uint32_t*>(scratch + L.idx_b);
    }
    int rc;
    if (key_bytes == 4)
        rc = run_passes<uint32_t>((const uint32_t*)keys_in, (uint32_t*)keys_out, vin, vout, (uint32_t)n, shift0, bits,
                                  npass, scratch, L, s);
    else
        rc = run_passes<uint64_t>((const uint64_t*)keys_in, (uint64_t*)keys_out, vin, vout, (uint32_t)n, shift0, bits,
                                  npass, scratch, L, s);
    if (rc) return rc;
    if (indirect) return col_gather(vals_in, vout, vals_out, n, value_bytes, 4, s);
    return 0;
}

}  // namespace

COL_API size_t col_sort_scratch_bytes(uint64_t n, uint32_t key_bytes, uint32_t value_bytes) {
    if (key_bytes != 4 && key_bytes != 8) return 0;
    return sort_layout(n, key_bytes, value_bytes).total;
}

COL_API int col_sort_pairs(const void* keys_in, void* keys_out, const void* vals_in, void* vals_out, uint64_t n,
                           uint32_t key_bytes, uint32_t value_bytes, uint32_t end_bit, void* scratch, void* stream) {
    if ((key_bytes != 4 && key_bytes != 8) || !scratch || (n && (!keys_in || !keys_out))) return COL_ERR_BAD_ARG;
    if (end_bit == 0 || end_bit > key_bytes * 8) end_bit = key_bytes * 8;
    const uint32_t npass = (end_bit + RADIX_BITS - 1) / RADIX_BITS;
    return sort_impl(keys_in, keys_out, vals_in, vals_out, n, key_bytes, value_bytes, 0, RADIX_BITS, npass, scratch,
                     col_stream(stream));
}

COL_API int col_radix_pass(const void* keys_in, void* keys_out, const void* vals_in, void* vals_out, uint64_t n,
                           uint32_t key_bytes, uint32_t value_bytes, uint32_t radix_bits, uint32_t pass, void* scratch,
                           void* stream) {
    if ((key_bytes != 4 && key_bytes != 8) || !scratch || (n && (!keys_in || !keys_out))) return COL_ERR_BAD_ARG;
    if (radix_bits == 0 || radix_bits > RADIX_BITS || (pass + 1) * radix_bits > key_bytes * 8) return COL_ERR_BAD_ARG;
    return sort_impl(keys_in, keys_out, vals_in, vals_out, n, key_bytes, value_bytes, pass * radix_bits, radix_bits, 1,
                     scratch, col_stream(stream));
}
