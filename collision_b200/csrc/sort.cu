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
                 uint32_t bits, uint32_t npass, const uint32_t* __restrict__ range) {
    __shared__ uint32_t sh[MAX_PASSES * RADIX];
    if (range) n = range[1];   // {offset, count} on the device: the host sized the launch for an upper bound
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

// RANGE (the sharded sort, col_shard_sort): the number of keys is only known on the device --
// range = {offset, count} -- so the host launches a fixed grid and every block takes tiles by ticket
// until they run out; `place` (the last pass) shifts the output by `offset`, the place of the
// sorted run in the whole array.
template <typename K, bool HAS_VALUES, bool RANGE>
__global__ void __launch_bounds__(SORT_THREADS, COL_SORT_CTAS)
onesweep_kernel(const K* __restrict__ keys_in, K* __restrict__ keys_out, const uint32_t* __restrict__ vals_in,
                uint32_t* __restrict__ vals_out, uint32_t n, const uint32_t* __restrict__ digit_base,
                uint32_t* __restrict__ status, unsigned* ticket, uint32_t shift, uint32_t mask, uint32_t bits,
                const uint32_t* __restrict__ range, uint32_t place) {
    constexpr int ITEMS = Tile<K>::ITEMS;
    constexpr int TILE = SORT_THREADS * ITEMS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_digit_start[RADIX];
    __shared__ uint32_t s_global_off[RADIX];
    __shared__ uint32_t s_scan_warp[RADIX / 32];
    __shared__ unsigned s_tile;

    uint32_t* s_whist = reinterpret_cast<uint32_t*>(smem_raw + sizeof(K) * skewed(TILE));
    if (RANGE) {
        n = range[1];
        if (place) {
            keys_out += range[0];
            if (HAS_VALUES) vals_out += range[0];
        }
    }
    do {
        if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
        for (int i = threadIdx.x; i < SORT_WARPS * RADIX; i += SORT_THREADS) s_whist[i] = 0;
        __syncthreads();
        const unsigned tile = s_tile;
        if (RANGE && (uint64_t)tile * TILE >= n) return;
        if ((uint64_t)(tile + 1) * TILE <= n)
            onesweep_tile<K, ITEMS, true, HAS_VALUES>(keys_in, keys_out, vals_in, vals_out, n, digit_base, status, shift,
                                                      mask, bits, tile, smem_raw, s_digit_start, s_global_off, s_scan_warp);
        else
            onesweep_tile<K, ITEMS, false, HAS_VALUES>(keys_in, keys_out, vals_in, vals_out, n, digit_base, status, shift,
                                                       mask, bits, tile, smem_raw, s_digit_start, s_global_off, s_scan_warp);
        if (RANGE) __syncthreads();   // the tile's shared memory is reused by the next one
    } while (RANGE);
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
               uint32_t shift0, uint32_t bits, uint32_t npass, char* scratch, const SortLayout& L, cudaStream_t s,
               const uint32_t* range = nullptr) {
    // range (device {offset, count}, see onesweep_kernel): n is then an upper bound, every buffer is read and
    // written from its start, except that the last pass writes at keys_out / vals_out + offset
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
        histogram_kernel<K><<<blocks, HIST_THREADS, 0, s>>>(keys_in, n, ghist, shift0, bits, npass, range);
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
    const bool col = sizeof(K) == 4 && column && !range && aligned32(keys_in) && aligned32(keys_out) && aligned32(tmp_keys) &&
                     (!has_vals || (aligned32(vals_in) && aligned32(vals_out) && aligned32(tmp_vals)));
    for (uint32_t p = 0; p < npass; p++) {
        // the last pass lands in *_out; earlier ones alternate tmp/out backwards from there
        const bool to_out = ((npass - 1 - p) & 1u) == 0;
        const K* kin = p == 0 ? keys_in : (to_out ? tmp_keys : keys_out);
        K* kout = to_out ? keys_out : tmp_keys;
        const uint32_t* vin = p == 0 ? vals_in : (to_out ? tmp_vals : vals_out);
        uint32_t* vout = to_out ? vals_out : tmp_vals;
        if (range) {
            // the input (the collected keys: scratch of the caller, writable) and tmp take turns, so that
            // *_out is written once, by the last pass, and only inside [offset, offset + count)
            const bool even = (p & 1u) == 0, last = p + 1 == npass;
            kin = even ? keys_in : tmp_keys;
            vin = even ? vals_in : tmp_vals;
            kout = last ? keys_out : (even ? tmp_keys : const_cast<K*>(keys_in));
            vout = last ? vals_out : (even ? tmp_vals : const_cast<uint32_t*>(vals_in));
        }
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
        if (range) {   // (key/value pairs only)
            constexpr size_t smem = onesweep_smem<K, true>();
            COL_CUDA(cudaFuncSetAttribute(onesweep_kernel<K, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            const unsigned resident = COL_NUM_SMS * COL_SORT_CTAS;
            const unsigned grid = L.tiles < resident ? (unsigned)L.tiles : resident;
            onesweep_kernel<K, true, true><<<grid, SORT_THREADS, smem, s>>>(
                kin, kout, vin, vout, n, ghist + p * RADIX, status, tickets + p, shift0 + p * bits, mask, bits, range,
                p + 1 == npass ? 1u : 0u);
        } else if (has_vals) {
            constexpr size_t smem = onesweep_smem<K, true>();
            COL_CUDA(cudaFuncSetAttribute(onesweep_kernel<K, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            onesweep_kernel<K, true, false><<<(unsigned)L.tiles, SORT_THREADS, smem, s>>>(
                kin, kout, vin, vout, n, ghist + p * RADIX, status, tickets + p, shift0 + p * bits, mask, bits, nullptr, 0u);
        } else {
            constexpr size_t smem = onesweep_smem<K, false>();
            COL_CUDA(cudaFuncSetAttribute(onesweep_kernel<K, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            onesweep_kernel<K, false, false><<<(unsigned)L.tiles, SORT_THREADS, smem, s>>>(
                kin, kout, nullptr, nullptr, n, ghist + p * RADIX, status, tickets + p, shift0 + p * bits, mask, bits, nullptr, 0u);
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

// ================================================================ sharded sort (multi-GPU)
// col_shard_sort: the part [pos_lo, pos_hi] of the stably sorted order of ALL n codes, without
// sorting the rest (collision_b200/sharded.py, mode "sharded_build": a rank only needs the sorted
// leaves of its own windows).  Every rank holds all n codes (4 B per sphere, all-gathered) and
//   1. counts their TOP digit (bits [end_bit - 8, end_bit)): the 256 bucket starts in the sorted order;
//   2. picks the buckets [d_lo, d_hi] that cover positions pos_lo..pos_hi: they hold `count` codes and
//      start at position `offset` (range_out = {offset, count, d_lo, d_hi}, all on the device -- no
//      host round trip, the step stays capturable);
//   3. collects (code, index) of exactly those codes, in input order (stable compaction: tiles in
//      ticket order, one 64-bit look-back word per tile);
//   4. sorts the collected pairs on all end_bit bits; the last pass writes them at
//      codes_sorted / ids_sorted + offset, their place in the order of the whole scene.
// Everything before `offset` belongs to smaller buckets, everything after to larger ones, and the
// LSD passes are stable, so positions [offset, offset + count) come out exactly as the full sort
// writes them.  Traffic: 8 B per code (count + collect) + 68 B per collected pair.  A scene that
// sits in one bucket degrades to the full sort (still correct).
namespace {

constexpr int COLLECT_THREADS = 256;
constexpr int COLLECT_VECS = 8;
constexpr int COLLECT_TILE = COLLECT_THREADS * COLLECT_VECS * 4;   // 8192 codes
constexpr int COLLECT_WARP_ITEMS = 32 * COLLECT_VECS * 4;

// starts = exclusive bucket starts (256); range_out = {offset, count, d_lo, d_hi}
__global__ void __launch_bounds__(RADIX) select_buckets_kernel(const uint32_t* __restrict__ starts, uint32_t n,
                                                               uint32_t pos_lo, uint32_t pos_hi,
                                                               uint32_t* __restrict__ range_out) {
    __shared__ uint32_t s_lo, s_hi;
    if (threadIdx.x == 0) { s_lo = 0; s_hi = 0; }
    __syncthreads();
    // the bucket that holds position p = the last one that starts at or before p (empty buckets share
    // their start with the next one, so the last such bucket is never empty for p < n)
    const uint32_t d = threadIdx.x, st = starts[d];
    if (st <= pos_lo) atomicMax(&s_lo, d);
    if (st <= pos_hi) atomicMax(&s_hi, d);
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t off = starts[s_lo], end = s_hi + 1 < RADIX ? starts[s_hi + 1] : n;
        range_out[0] = off;
        range_out[1] = end - off;
        range_out[2] = s_lo;
        range_out[3] = s_hi;
    }
}

__global__ void __launch_bounds__(COLLECT_THREADS)
collect_kernel(const uint32_t* __restrict__ codes, uint32_t n, uint32_t shift, uint32_t mask, const uint32_t* __restrict__ range,
               uint32_t* __restrict__ keys_out, uint32_t* __restrict__ ids_out, unsigned* ticket, uint64_t* status) {
    __shared__ unsigned s_tile;
    __shared__ uint32_t s_warp_total[COLLECT_THREADS / 32];
    __shared__ uint32_t s_tile_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const unsigned tile = s_tile;
    const uint32_t tile_base = tile * (uint32_t)COLLECT_TILE;
    const unsigned warp = threadIdx.x >> 5, lane = lane_id();
    const uint32_t warp_base = tile_base + warp * COLLECT_WARP_ITEMS;
    const bool full = (uint64_t)tile_base + COLLECT_TILE <= n;
    const uint32_t d_lo = range[2], span = range[3] - d_lo;

    uint4 v[COLLECT_VECS];
    uint32_t sel[COLLECT_VECS];     // 4 bits: which of the four codes of the vector are collected
    uint32_t excl[COLLECT_VECS];    // collected before mine inside the warp's part of the tile
    uint32_t row_base = 0;
#pragma unroll
    for (int k = 0; k < COLLECT_VECS; k++) {
        const uint32_t e = warp_base + (uint32_t)(k * 32 + lane) * 4;
        if (full) {
            v[k] = ld_stream_u4(reinterpret_cast<const uint4*>(codes + e));
        } else {
            v[k].x = e + 0 < n ? codes[e + 0] : 0u;
            v[k].y = e + 1 < n ? codes[e + 1] : 0u;
            v[k].z = e + 2 < n ? codes[e + 2] : 0u;
            v[k].w = e + 3 < n ? codes[e + 3] : 0u;
        }
    }
#pragma unroll
    for (int k = 0; k < COLLECT_VECS; k++) {
        const uint32_t e = warp_base + (uint32_t)(k * 32 + lane) * 4;
        uint32_t m = 0;
        m |= ((((v[k].x >> shift) & mask) - d_lo) <= span && (full || e + 0 < n)) ? 1u : 0u;
        m |= ((((v[k].y >> shift) & mask) - d_lo) <= span && (full || e + 1 < n)) ? 2u : 0u;
        m |= ((((v[k].z >> shift) & mask) - d_lo) <= span && (full || e + 2 < n)) ? 4u : 0u;
        m |= ((((v[k].w >> shift) & mask) - d_lo) <= span && (full || e + 3 < n)) ? 8u : 0u;
        sel[k] = m;
        const uint32_t c = (uint32_t)__popc(m);
        uint32_t inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += up;
        }
        excl[k] = row_base + inc - c;
        row_base += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) s_warp_total[warp] = row_base;
    __syncthreads();
    uint32_t warp_prefix = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < COLLECT_THREADS / 32; w++) {
        const uint32_t t = s_warp_total[w];
        if (w < (int)warp) warp_prefix += t;
        tile_total += t;
    }
    // decoupled look-back by warp 0 over {flag:2, count:62} words, 32 predecessors per step
    if (warp == 0) {
        constexpr uint64_t PARTIAL = 1ull << 62, INCLUSIVE = 2ull << 62, VALUE = (1ull << 62) - 1;
        uint64_t prefix = 0;
        if (tile == 0) {
            if (lane == 0) st_relaxed_u64(status, INCLUSIVE | tile_total);
        } else {
            if (lane == 0) st_relaxed_u64(status + tile, PARTIAL | tile_total);
            int64_t look = (int64_t)tile - 1 - lane;
            while (true) {
                uint64_t st = INCLUSIVE;
                if (look >= 0) {
                    do { st = ld_relaxed_u64(status + look); } while ((st >> 62) == 0);
                }
                const unsigned inc_mask = __ballot_sync(0xffffffffu, (st >> 62) == 2);
                const unsigned first = inc_mask ? (unsigned)__ffs(inc_mask) - 1 : 31u;
                uint64_t c = (lane <= first) ? (st & VALUE) : 0;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                prefix += c;
                if (inc_mask) break;
                look -= 32;
            }
            if (lane == 0) st_relaxed_u64(status + tile, INCLUSIVE | (prefix + tile_total));
        }
        if (lane == 0) s_tile_prefix = (uint32_t)prefix;
    }
    __syncthreads();
    const uint32_t base = s_tile_prefix + warp_prefix;
#pragma unroll
    for (int k = 0; k < COLLECT_VECS; k++) {
        const uint32_t e = warp_base + (uint32_t)(k * 32 + lane) * 4;
        uint32_t dst = base + excl[k];
        const uint32_t m = sel[k];
#ifdef COL_DEBUG
        if (m && dst + __popc(m) > range[1]) __trap();   // more collected than the buckets hold
#endif
        if (m & 1u) { keys_out[dst] = v[k].x; ids_out[dst] = e + 0; dst++; }
        if (m & 2u) { keys_out[dst] = v[k].y; ids_out[dst] = e + 1; dst++; }
        if (m & 4u) { keys_out[dst] = v[k].z; ids_out[dst] = e + 2; dst++; }
        if (m & 8u) { keys_out[dst] = v[k].w; ids_out[dst] = e + 3; dst++; }
    }
}

struct ShardSortLayout {
    size_t hist, ticket, status, keys, vals, sort, total;
    uint64_t tiles;
};

ShardSortLayout shard_sort_layout(uint64_t n) {
    ShardSortLayout L;
    L.tiles = (n + COLLECT_TILE - 1) / COLLECT_TILE;
    if (L.tiles == 0) L.tiles = 1;
    size_t off = 0;
    L.hist = off; off += align_up(sizeof(uint32_t) * RADIX);
    L.ticket = off; off += align_up(sizeof(uint32_t));
    L.status = off; off += align_up(sizeof(uint64_t) * L.tiles);
    L.keys = off; off += align_up((size_t)n * 4);
    L.vals = off; off += align_up((size_t)n * 4);
    L.sort = off; off += sort_layout(n, 4, 4).total;
    L.total = off;
    return L;
}

}  // namespace

COL_API size_t col_shard_sort_scratch_bytes(uint64_t n) { return shard_sort_layout(n).total; }

COL_API int col_shard_sort(const uint32_t* codes, uint32_t n, uint32_t pos_lo, uint32_t pos_hi, uint32_t end_bit,
                           uint32_t* codes_sorted, uint32_t* ids_sorted, uint32_t* range_out, void* scratch_v,
                           void* stream) {
    if (!codes || !codes_sorted || !ids_sorted || !range_out || !scratch_v) return COL_ERR_BAD_ARG;
    if (n == 0 || pos_lo > pos_hi || pos_lo >= n) return COL_ERR_BAD_ARG;
    if (n > (1u << 30)) return COL_ERR_TOO_LARGE;
    if (end_bit == 0 || end_bit > 32) end_bit = 32;
    if (pos_hi >= n) pos_hi = n - 1;
    if ((reinterpret_cast<uintptr_t>(codes) & 15u) != 0) return COL_ERR_BAD_ARG;   // 128-bit loads
    cudaStream_t s = col_stream(stream);
    char* scratch = reinterpret_cast<char*>(scratch_v);
    const ShardSortLayout L = shard_sort_layout(n);
    uint32_t* hist = reinterpret_cast<uint32_t*>(scratch + L.hist);
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch + L.ticket);
    uint64_t* status = reinterpret_cast<uint64_t*>(scratch + L.status);
    uint32_t* keys = reinterpret_cast<uint32_t*>(scratch + L.keys);
    uint32_t* vals = reinterpret_cast<uint32_t*>(scratch + L.vals);
    const uint32_t top_shift = end_bit > RADIX_BITS ? end_bit - RADIX_BITS : 0;
    const uint32_t top_bits = end_bit > RADIX_BITS ? RADIX_BITS : end_bit;
    COL_CUDA(cudaMemsetAsync(scratch, 0, L.keys, s));   // histogram, ticket, look-back words
    {
        const uint64_t want = ((uint64_t)n + HIST_THREADS * HIST_ITEMS - 1) / (HIST_THREADS * HIST_ITEMS);
        const unsigned blocks = (unsigned)(want > COL_NUM_SMS * 4 ? COL_NUM_SMS * 4 : want);
        histogram_kernel<uint32_t><<<blocks, HIST_THREADS, 0, s>>>(codes, n, hist, top_shift, top_bits, 1, nullptr);
        COL_CHECK_LAUNCH();
        digit_base_kernel<<<1, RADIX, 0, s>>>(hist);
        COL_CHECK_LAUNCH();
    }
    select_buckets_kernel<<<1, RADIX, 0, s>>>(hist, n, pos_lo, pos_hi, range_out);
    COL_CHECK_LAUNCH();
    collect_kernel<<<(unsigned)L.tiles, COLLECT_THREADS, 0, s>>>(codes, n, top_shift, (1u << top_bits) - 1, range_out, keys, vals,
                                                                ticket, status);
    COL_CHECK_LAUNCH();
    const uint32_t npass = (end_bit + RADIX_BITS - 1) / RADIX_BITS;
    const SortLayout SL = sort_layout(n, 4, 4);
    return run_passes<uint32_t>(keys, codes_sorted, vals, ids_sorted, n, 0, RADIX_BITS, npass, scratch + L.sort, SL, s,
                                range_out);
}

// ================================================================ sharded sort, exchanged (multi-GPU)
// The same selection as col_shard_sort, but without ANY rank ever holding all the codes or the whole
// scene: the spheres a rank's windows need are PUSHED to it by their holders, straight into its HBM
// over NVLink (peer stores into buffers mapped with CUDA IPC), compacted on the way.
//   col_shard_count   : top-digit counts of the rank's OWN slice of the codes (256 words; the caller
//                       all-gathers them: 1 KB per rank instead of 4 B per sphere);
//   col_shard_push    : from the gathered counts every rank derives the same plan -- bucket starts, and for
//                       every rank r the buckets covering its positions -- plus where ITS contribution
//                       to r's receive buffers starts (the spheres of lower ranks' slices in those buckets
//                       come first: input order).  Then one stable compaction of the own slice per
//                       destination, writing {code, sphere id, coordinates, radius} to the destination;
//   (the caller runs a collective here: all pushes have landed once every rank has passed it)
//   col_shard_sort_received : stable sort of the received codes (value = place in the receive buffers),
//                       placed at the rank's positions of the global order; ids_sorted = the received
//                       ids in that order.  The build reads coordinates and radii through the sorted
//                       places, from the receive buffers.
// The received list is in input order (rank-major, then slice order), exactly what col_shard_sort's
// collect pass produces from all the codes: positions come out bit-identical to the full sort.
namespace {

constexpr uint32_t PLAN_WORDS = 8;    // per destination: {offset, count, first bucket, last bucket, my base, -, -, -}
constexpr uint32_t NO_BUCKET = 0x100; // first = last = NO_BUCKET: the destination takes nothing

// hist_all: world x 256 counts (rank-major); pos: world x {pos_lo, pos_hi} (pos_lo > pos_hi: nothing)
__global__ void __launch_bounds__(RADIX) plan_kernel(const uint32_t* __restrict__ hist_all, uint32_t world, uint32_t me,
                                                     const uint32_t* __restrict__ pos, uint32_t n,
                                                     uint32_t* __restrict__ plan) {
    __shared__ uint32_t s_start[RADIX + 1];
    __shared__ uint32_t s_warp[RADIX / 32];
    __shared__ uint32_t s_lo, s_hi, s_base;
    const uint32_t d = threadIdx.x;
    uint32_t total = 0, before_me = 0;
    for (uint32_t r = 0; r < world; r++) {
        const uint32_t c = hist_all[r * RADIX + d];
        total += c;
        if (r < me) before_me += c;
    }
    uint32_t inc = total;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane_id() >= o) inc += up;
    }
    if (lane_id() == 31) s_warp[d >> 5] = inc;
    __syncthreads();
    uint32_t base = 0;
    for (uint32_t w = 0; w < (d >> 5); w++) base += s_warp[w];
    s_start[d] = base + inc - total;
    if (d == RADIX - 1) s_start[RADIX] = base + inc;   // = n
    __syncthreads();
    for (uint32_t r = 0; r < world; r++) {
        const uint32_t pos_lo = pos[2 * r], pos_hi = min(pos[2 * r + 1], n - 1);
        if (d == 0) { s_lo = 0; s_hi = 0; s_base = 0; }
        __syncthreads();
        const bool any = pos_lo <= pos_hi && pos_lo < n;
        if (any) {
            if (s_start[d] <= pos_lo) atomicMax(&s_lo, d);
            if (s_start[d] <= pos_hi) atomicMax(&s_hi, d);
        }
        __syncthreads();
        if (any && d >= s_lo && d <= s_hi && before_me) atomicAdd(&s_base, before_me);
        __syncthreads();
        if (d == 0) {
            uint32_t* p = plan + r * PLAN_WORDS;
            p[0] = any ? s_start[s_lo] : 0u;
            p[1] = any ? s_start[s_hi + 1] - s_start[s_lo] : 0u;
            p[2] = any ? s_lo : NO_BUCKET;
            p[3] = any ? s_hi : NO_BUCKET;
            p[4] = s_base;
            p[5] = p[6] = p[7] = 0u;
        }
        __syncthreads();
    }
}

// collect_kernel with the scene riding along and the output somewhere else: entries go to
// dst_* + plan[4] + (their rank among the collected entries of this slice).  W = words per scalar.
template <int W>
__global__ void __launch_bounds__(COLLECT_THREADS)
collect_push_kernel(const uint32_t* __restrict__ codes, uint32_t n, uint32_t id_base, uint32_t shift, uint32_t mask,
                    const uint32_t* __restrict__ plan, const uint4* __restrict__ coords, const uint32_t* __restrict__ radii,
                    uint32_t* __restrict__ dst_codes, uint32_t* __restrict__ dst_ids, uint4* __restrict__ dst_coords,
                    uint32_t* __restrict__ dst_radii, unsigned* ticket, uint64_t* status) {
    __shared__ unsigned s_tile;
    __shared__ uint32_t s_warp_total[COLLECT_THREADS / 32];
    __shared__ uint32_t s_tile_prefix;
    __shared__ unsigned short s_off[COLLECT_TILE];
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const unsigned tile = s_tile;
    const uint32_t tile_base = tile * (uint32_t)COLLECT_TILE;
    const unsigned warp = threadIdx.x >> 5, lane = lane_id();
    const uint32_t warp_base = tile_base + warp * COLLECT_WARP_ITEMS;
    const bool full = (uint64_t)tile_base + COLLECT_TILE <= n;
    const uint32_t d_lo = plan[2], span = plan[3] - d_lo, out0 = plan[4];

    uint4 v[COLLECT_VECS];
    uint32_t sel[COLLECT_VECS], excl[COLLECT_VECS];
    uint32_t row_base = 0;
#pragma unroll
    for (int k = 0; k < COLLECT_VECS; k++) {
        const uint32_t e = warp_base + (uint32_t)(k * 32 + lane) * 4;
        if (full) {
            v[k] = *reinterpret_cast<const uint4*>(codes + e);
        } else {
            v[k].x = e + 0 < n ? codes[e + 0] : 0u;
            v[k].y = e + 1 < n ? codes[e + 1] : 0u;
            v[k].z = e + 2 < n ? codes[e + 2] : 0u;
            v[k].w = e + 3 < n ? codes[e + 3] : 0u;
        }
    }
#pragma unroll
    for (int k = 0; k < COLLECT_VECS; k++) {
        const uint32_t e = warp_base + (uint32_t)(k * 32 + lane) * 4;
        uint32_t m = 0;
        m |= ((((v[k].x >> shift) & mask) - d_lo) <= span && (full || e + 0 < n)) ? 1u : 0u;
        m |= ((((v[k].y >> shift) & mask) - d_lo) <= span && (full || e + 1 < n)) ? 2u : 0u;
        m |= ((((v[k].z >> shift) & mask) - d_lo) <= span && (full || e + 2 < n)) ? 4u : 0u;
        m |= ((((v[k].w >> shift) & mask) - d_lo) <= span && (full || e + 3 < n)) ? 8u : 0u;
        sel[k] = m;
        const uint32_t c = (uint32_t)__popc(m);
        uint32_t inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += up;
        }
        excl[k] = row_base + inc - c;
        row_base += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) s_warp_total[warp] = row_base;
    __syncthreads();
    uint32_t warp_prefix = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < COLLECT_THREADS / 32; w++) {
        const uint32_t t = s_warp_total[w];
        if (w < (int)warp) warp_prefix += t;
        tile_total += t;
    }
    if (warp == 0) {
        constexpr uint64_t PARTIAL = 1ull << 62, INCLUSIVE = 2ull << 62, VALUE = (1ull << 62) - 1;
        uint64_t prefix = 0;
        if (tile == 0) {
            if (lane == 0) st_relaxed_u64(status, INCLUSIVE | tile_total);
        } else {
            if (lane == 0) st_relaxed_u64(status + tile, PARTIAL | tile_total);
            int64_t look = (int64_t)tile - 1 - lane;
            while (true) {
                uint64_t st = INCLUSIVE;
                if (look >= 0) {
                    do { st = ld_relaxed_u64(status + look); } while ((st >> 62) == 0);
                }
                const unsigned inc_mask = __ballot_sync(0xffffffffu, (st >> 62) == 2);
                const unsigned first = inc_mask ? (unsigned)__ffs(inc_mask) - 1 : 31u;
                uint64_t c = (lane <= first) ? (st & VALUE) : 0;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                prefix += c;
                if (inc_mask) break;
                look -= 32;
            }
            if (lane == 0) st_relaxed_u64(status + tile, INCLUSIVE | (prefix + tile_total));
        }
        if (lane == 0) s_tile_prefix = (uint32_t)prefix;
    }
    __syncthreads();
    // the tile's collected entries, in order, as offsets into the tile: consecutive threads then write
    // consecutive entries, so that what crosses NVLink are full lines, not 4- and 16-byte pieces
#pragma unroll
    for (int k = 0; k < COLLECT_VECS; k++) {
        const uint32_t e = warp * COLLECT_WARP_ITEMS + (uint32_t)(k * 32 + lane) * 4;
        uint32_t p = warp_prefix + excl[k];
        const uint32_t m = sel[k];
#pragma unroll
        for (int h = 0; h < 4; h++)
            if (m & (1u << h)) s_off[p++] = (unsigned short)(e + h);
    }
    __syncthreads();
    const uint32_t base = out0 + s_tile_prefix;
#ifdef COL_DEBUG
    if (tile_total && base + tile_total > plan[1]) __trap();   // more than the destination's buckets hold
#endif
    for (uint32_t p = threadIdx.x; p < tile_total; p += COLLECT_THREADS) {
        const uint32_t src = tile_base + s_off[p], dst = base + p;
        dst_codes[dst] = __ldg(codes + src);
        dst_ids[dst] = id_base + src;
#pragma unroll
        for (int w = 0; w < W; w++) {
            dst_coords[(size_t)dst * W + w] = ld_stream_u4(coords + (size_t)src * W + w);
            dst_radii[(size_t)dst * W + w] = ld_stream_u32(radii + (size_t)src * W + w);
        }
    }
}

__global__ void __launch_bounds__(256) iota_range_kernel(uint32_t* p, const uint32_t* __restrict__ range) {
    const uint32_t n = range[1];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = i;
}

// ids_sorted[offset + i] = recv_ids[place_sorted[offset + i]]
__global__ void __launch_bounds__(256) gather_ids_range_kernel(uint32_t* __restrict__ ids_sorted,
                                                               const uint32_t* __restrict__ place_sorted,
                                                               const uint32_t* __restrict__ recv_ids,
                                                               const uint32_t* __restrict__ range) {
    const uint32_t off = range[0], n = range[1];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        ids_sorted[off + i] = __ldg(recv_ids + place_sorted[off + i]);
}

struct PushLayout {
    size_t tickets, status, status_stride, total;
    uint64_t tiles;
};

PushLayout push_layout(uint64_t n_own, uint32_t world) {
    PushLayout L;
    L.tiles = (n_own + COLLECT_TILE - 1) / COLLECT_TILE;
    if (L.tiles == 0) L.tiles = 1;
    size_t off = 0;
    L.tickets = off; off += align_up(sizeof(uint32_t) * world);
    L.status = off;
    L.status_stride = align_up(sizeof(uint64_t) * L.tiles);
    off += L.status_stride * world;
    L.total = off;
    return L;
}

}  // namespace

COL_API int col_shard_count(const uint32_t* codes_own, uint32_t n_own, uint32_t end_bit, uint32_t* hist256, void* stream) {
    if (!hist256 || (n_own && !codes_own)) return COL_ERR_BAD_ARG;
    if (end_bit == 0 || end_bit > 32) end_bit = 32;
    cudaStream_t s = col_stream(stream);
    COL_CUDA(cudaMemsetAsync(hist256, 0, sizeof(uint32_t) * RADIX, s));
    if (n_own == 0) return 0;
    const uint32_t top_shift = end_bit > RADIX_BITS ? end_bit - RADIX_BITS : 0;
    const uint32_t top_bits = end_bit > RADIX_BITS ? RADIX_BITS : end_bit;
    const uint64_t want = ((uint64_t)n_own + HIST_THREADS * HIST_ITEMS - 1) / (HIST_THREADS * HIST_ITEMS);
    const unsigned blocks = (unsigned)(want > COL_NUM_SMS * 4 ? COL_NUM_SMS * 4 : want);
    histogram_kernel<uint32_t><<<blocks, HIST_THREADS, 0, s>>>(codes_own, n_own, hist256, top_shift, top_bits, 1, nullptr);
    COL_CHECK_LAUNCH();
    return 0;
}

COL_API size_t col_shard_push_scratch_bytes(uint64_t n_own, uint32_t world) { return push_layout(n_own, world).total; }

COL_API int col_shard_push(const uint32_t* hist_all, uint32_t world, uint32_t my_rank, const uint32_t* pos_table,
                           uint32_t n, const uint32_t* codes_own, uint32_t n_own, uint32_t id_base,
                           const void* coords_own, const void* radii_own, int dtype, uint32_t end_bit,
                           const uint64_t* h_recv_codes, const uint64_t* h_recv_ids, const uint64_t* h_recv_coords,
                           const uint64_t* h_recv_radii, uint32_t* plan, void* scratch_v, void* stream) {
    if (!hist_all || !pos_table || !plan || !scratch_v || !h_recv_codes || !h_recv_ids || !h_recv_coords || !h_recv_radii ||
        world == 0 || my_rank >= world || n == 0)
        return COL_ERR_BAD_ARG;
    if (n_own && (!codes_own || !coords_own || !radii_own)) return COL_ERR_BAD_ARG;
    if (dtype != COL_F32 && dtype != COL_F64) return COL_ERR_UNSUPPORTED;
    if (n > (1u << 30)) return COL_ERR_TOO_LARGE;
    if ((reinterpret_cast<uintptr_t>(codes_own) & 15u) != 0 || (reinterpret_cast<uintptr_t>(coords_own) & 15u) != 0)
        return COL_ERR_BAD_ARG;
    if (end_bit == 0 || end_bit > 32) end_bit = 32;
    cudaStream_t s = col_stream(stream);
    char* scratch = reinterpret_cast<char*>(scratch_v);
    const PushLayout L = push_layout(n_own, world);
    const uint32_t top_shift = end_bit > RADIX_BITS ? end_bit - RADIX_BITS : 0;
    const uint32_t top_bits = end_bit > RADIX_BITS ? RADIX_BITS : end_bit;
    plan_kernel<<<1, RADIX, 0, s>>>(hist_all, world, my_rank, pos_table, n, plan);
    COL_CHECK_LAUNCH();
    if (n_own == 0) return 0;
    COL_CUDA(cudaMemsetAsync(scratch, 0, L.total, s));
    for (uint32_t k = 0; k < world; k++) {
        const uint32_t r = (my_rank + k) % world;   // every rank starts with its own buffers, then round the ring
        unsigned* ticket = reinterpret_cast<unsigned*>(scratch + L.tickets) + r;
        uint64_t* status = reinterpret_cast<uint64_t*>(scratch + L.status + L.status_stride * r);
        if (dtype == COL_F32)
            collect_push_kernel<1><<<(unsigned)L.tiles, COLLECT_THREADS, 0, s>>>(
                codes_own, n_own, id_base, top_shift, (1u << top_bits) - 1, plan + r * PLAN_WORDS, (const uint4*)coords_own,
                (const uint32_t*)radii_own, (uint32_t*)h_recv_codes[r], (uint32_t*)h_recv_ids[r], (uint4*)h_recv_coords[r],
                (uint32_t*)h_recv_radii[r], ticket, status);
        else
            collect_push_kernel<2><<<(unsigned)L.tiles, COLLECT_THREADS, 0, s>>>(
                codes_own, n_own, id_base, top_shift, (1u << top_bits) - 1, plan + r * PLAN_WORDS, (const uint4*)coords_own,
                (const uint32_t*)radii_own, (uint32_t*)h_recv_codes[r], (uint32_t*)h_recv_ids[r], (uint4*)h_recv_coords[r],
                (uint32_t*)h_recv_radii[r], ticket, status);
        COL_CHECK_LAUNCH();
    }
    return 0;
}

COL_API size_t col_shard_sort_received_scratch_bytes(uint64_t n) {
    return align_up((size_t)n * 4) + sort_layout(n, 4, 4).total;
}

COL_API int col_shard_sort_received(uint32_t* recv_codes, const uint32_t* recv_ids, uint32_t n, const uint32_t* range,
                                    uint32_t end_bit, uint32_t* codes_sorted, uint32_t* place_sorted, uint32_t* ids_sorted,
                                    void* scratch_v, void* stream) {
    if (!recv_codes || !recv_ids || !range || !codes_sorted || !place_sorted || !ids_sorted || !scratch_v || n == 0)
        return COL_ERR_BAD_ARG;
    if (n > (1u << 30)) return COL_ERR_TOO_LARGE;
    if (end_bit == 0 || end_bit > 32) end_bit = 32;
    cudaStream_t s = col_stream(stream);
    char* scratch = reinterpret_cast<char*>(scratch_v);
    uint32_t* places = reinterpret_cast<uint32_t*>(scratch);
    iota_range_kernel<<<COL_NUM_SMS * 8, 256, 0, s>>>(places, range);
    COL_CHECK_LAUNCH();
    const uint32_t npass = (end_bit + RADIX_BITS - 1) / RADIX_BITS;
    const SortLayout SL = sort_layout(n, 4, 4);
    const int rc = run_passes<uint32_t>(recv_codes, codes_sorted, places, place_sorted, n, 0, RADIX_BITS, npass,
                                        scratch + align_up((size_t)n * 4), SL, s, range);
    if (rc) return rc;
    gather_ids_range_kernel<<<COL_NUM_SMS * 8, 256, 0, s>>>(ids_sorted, place_sorted, recv_ids, range);
    COL_CHECK_LAUNCH();
    return 0;
}
