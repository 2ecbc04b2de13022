// sort_col.cuh -- the one-sweep pass for 32-bit keys, "column counter" form.
//
// Same contract as onesweep_kernel in sort.cu (one launch per 8-bit digit, tiles in ticket
// order, decoupled look-back on {flag:2, count:30} words, stable), different tile algorithm.
// The ballot ranking of sort.cu costs ~67 warp instructions per row of 32 keys, most of
// them on the integer-ALU pipe (one per 2 cycles per scheduler on B200), which is what
// bounds it at ~45 % of the HBM roofline.  Here a tile is ranked by COUNTING, with plain
// byte loads/stores and packed 32-bit adds -- no votes, no popc, no atomics:
//
//   * blocked arrangement: thread t owns keys [32t, 32t+32) of the 8192-key tile (four
//     256-bit loads), so tile order = (thread, item) and a thread's keys are ranked in
//     program order;
//   * a byte counter per (digit, thread): 256 x 256 B = 64 KB of shared memory.  The bytes of
//     threads t, t+64, t+128, t+192 share one word, word (d, x) = d * 64 + x with x = t & 63: the
//     32 lanes of a warp always hit 32 different banks whatever their digits are;
//   * scan A: every 8 consecutive threads (one word column of 8 words, all four bytes at
//     once) get their exclusive prefix by packed adds -- at most 7 x 32 = 224 per byte, so no
//     byte ever carries -- and the group's four totals go to a 16-bit table;
//   * scan B: one thread per digit turns the 32 group totals of its digit into run starts
//     inside the sorted tile (two packed 16-bit running sums), after a block scan over the
//     digit totals, which are published for the look-back as soon as they are known;
//   * ranking: rank = table[digit][group] + scanned counter[digit][thread] + the key's occurrence
//     index inside its thread (remembered from the count) -- one byte load and one 16-bit load per key;
//   * the sorted tile (keys, or key/value pairs as 8-byte elements) re-uses the counters'
//     64 KB, XOR-swizzled (index ^ (index >> 5)) instead of padded, and goes out in runs of
//     consecutive addresses exactly like sort.cu.
//
// tools/sort_col_model.py is a numpy model of the tile algorithm (same layouts, same packed
// arithmetic incl. the all-keys-in-one-bin corner where a group total is 256).
#pragma once
#include "common.cuh"

namespace colsort {

constexpr int T = 256;             // threads per block
constexpr int I = 32;              // keys per thread
constexpr int TILE = T * I;        // 8192 keys
constexpr int NDIG = 256;
constexpr int CNT_BYTES = NDIG * T;          // 65536: byte counter per (digit, thread)
constexpr int BASE_ROW = 80;                 // bytes per digit row of the 16-bit table (64 used; 80 keeps rows on distinct banks)
constexpr int BASE_BYTES = NDIG * BASE_ROW;  // 20480

constexpr uint32_t ST_PARTIAL = 1u << 30;
constexpr uint32_t ST_INCLUSIVE = 2u << 30;
constexpr uint32_t ST_VALUE = (1u << 30) - 1;

template <bool HAS_VALUES> constexpr size_t smem_bytes() { return CNT_BYTES + BASE_BYTES; }

// digit of a key; BYTE >= 0: the digit is byte BYTE of the key (one PRMT), else (k >> shift) & mask
template <int BYTE> __device__ __forceinline__ uint32_t digit_of(uint32_t k, uint32_t shift, uint32_t mask) {
    if (BYTE >= 0) return __byte_perm(k, 0, 0x4440 | BYTE);
    return (k >> shift) & mask;
}
// byte offset of counter (digit of k, this thread): digit * 256 + toff, toff = (t & 63) * 4 + (t >> 6) < 256
template <int BYTE> __device__ __forceinline__ uint32_t cnt_addr(uint32_t k, uint32_t toff, uint32_t shift, uint32_t mask) {
    if (BYTE >= 0) return __byte_perm(k, toff, 0x5504 | (BYTE << 4));   // {toff.b0, k.b[BYTE], 0, 0}
    return (((k >> shift) & mask) << 8) | toff;
}

__device__ __forceinline__ void ld_v8(const uint32_t* p, uint32_t* r) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "l"(p));
}

template <int BYTE, bool FULL, bool HAS_VALUES>
__device__ __forceinline__ void col_tile(const uint32_t* __restrict__ keys_in, uint32_t* __restrict__ keys_out,
                                         const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out,
                                         uint32_t n, const uint32_t* __restrict__ digit_base,
                                         uint32_t* __restrict__ status, uint32_t shift, uint32_t mask, unsigned tile,
                                         unsigned char* smem, uint32_t* s_goff, uint32_t* s_dstart, uint32_t* s_warp) {
    unsigned char* cnt = smem;                       // counters, later the sorted tile
    unsigned char* base = smem + CNT_BYTES;          // 16-bit table, 4 x run start
    const unsigned t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const unsigned b = t >> 6, x = t & 63, g = x >> 3;
    // word x of a digit row sits at x ^ 4 for x >= 32: the two halves of groups 4..7 are swapped in
    // memory, so that the 128-bit accesses of scan A (8 consecutive groups per quarter warp, every
    // lane asking for the same logical half) fall on 8 different bank quads instead of 4
    const uint32_t toff = (x ^ ((x >> 5) << 2)) * 4 + b;
    const uint32_t tile_base = tile * (uint32_t)TILE;
    const uint32_t valid = FULL ? (uint32_t)TILE : n - tile_base;

    // ---- load: 32 consecutive keys per thread
    uint32_t key[I];
    {
        const uint32_t* src = keys_in + tile_base + t * I;
        if (FULL) {
#pragma unroll
            for (int v = 0; v < I / 8; v++) ld_v8(src + v * 8, key + v * 8);
        } else {
#pragma unroll
            for (int i = 0; i < I; i++) key[i] = (t * I + i) < valid ? src[i] : 0xFFFFFFFFu;   // padding sorts last
        }
    }

    // ---- count: one byte per (digit, thread); program order per thread, no atomics.  What the
    // counter held before a key's increment is the number of EARLIER keys of this thread with the
    // same digit: kept (a byte per key, four per register) so that the ranking needs no second
    // read-modify-write
    uint32_t occ[I / 4];
#pragma unroll
    for (int i = 0; i < I; i += 4) {
        uint32_t q4 = 0;
#pragma unroll
        for (int h = 0; h < 4; h++) {
            unsigned char* p = cnt + cnt_addr<BYTE>(key[i + h], toff, shift, mask);
            const uint32_t q = *p;
            *p = (unsigned char)(q + 1);
            q4 |= q << (8 * h);
        }
        occ[i >> 2] = q4;
    }
    __syncthreads();

    // ---- scan A: unit u = (digit u >> 3, group u & 7) = 8 words = the counters of 8 consecutive
    // threads x 4 byte lanes.  Exclusive prefix over the 8 words by packed adds (<= 224 per byte),
    // the group's four totals (<= 256 each, hence summed in two halves) to the 16-bit table.
    const unsigned flip = (t >> 2) & 1u;             // group (u & 7) >= 4: its halves are swapped (see toff)
#pragma unroll
    for (int k = 0; k < (NDIG * 8) / T; k++) {
        const unsigned u = t + T * k;
        uint4* p = reinterpret_cast<uint4*>(cnt + u * 32);
        uint4* pa = p + flip;                        // threads 0..3 of the group
        uint4* pc = p + (flip ^ 1u);                 // threads 4..7
        const uint4 a = *pa, c = *pc;
        const uint32_t e1 = a.x, e2 = e1 + a.y, e3 = e2 + a.z, e4 = e3 + a.w;
        const uint32_t e5 = e4 + c.x, e6 = e5 + c.y, e7 = e6 + c.z;
        const uint32_t thi = e7 - e4 + c.w;          // threads 4..7 of the group, per byte lane
        *pa = make_uint4(0u, e1, e2, e3);
        *pc = make_uint4(e4, e5, e6, e7);
        const uint32_t A = (e4 & 0x00ff00ffu) + (thi & 0x00ff00ffu);                    // lanes b0 | b2 << 16
        const uint32_t Bv = ((e4 >> 8) & 0x00ff00ffu) + ((thi >> 8) & 0x00ff00ffu);     // lanes b1 | b3 << 16
        *reinterpret_cast<uint2*>(base + (u >> 3) * BASE_ROW + (u & 7) * 8) = make_uint2(A, Bv);
    }
    __syncthreads();

    // ---- scan B: thread = digit.  Row = {A_g, B_g} for g = 0..7.  Tile order is thread order =
    // byte lane b major (b = t >> 6), then group g: stripe totals first, then two packed running sums.
    uint32_t real;      // real (non-padding) keys of digit t in this tile
    uint32_t ds;        // first slot of digit t in the sorted tile
    {
        uint4* row = reinterpret_cast<uint4*>(base + t * BASE_ROW);
        uint4 r0 = row[0], r1 = row[1], r2 = row[2], r3 = row[3];      // {A0 B0 A1 B1} {A2 B2 A3 B3} ...
        const uint32_t sumA = r0.x + r0.z + r1.x + r1.z + r2.x + r2.z + r3.x + r3.z;   // S0 | S2 << 16  (<= 2048 each)
        const uint32_t sumB = r0.y + r0.w + r1.y + r1.w + r2.y + r2.w + r3.y + r3.w;   // S1 | S3 << 16
        const uint32_t S0 = sumA & 0xffffu, S2 = sumA >> 16, S1 = sumB & 0xffffu, S3 = sumB >> 16;
        const uint32_t tot = S0 + S1 + S2 + S3;
        // the padding keys of a partial tile all carry the all-ones digit
        real = tot - ((!FULL && t == mask) ? (uint32_t)TILE - valid : 0u);
        st_relaxed_u32(status + (size_t)tile * NDIG + t, (tile == 0 ? ST_INCLUSIVE : ST_PARTIAL) | real);
        uint32_t inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= (unsigned)o) inc += up;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        ds = inc - tot;
        for (unsigned w = 0; w < warp; w++) ds += s_warp[w];
        s_dstart[t] = ds;
        // run starts x 4 (byte offsets of 4-byte slots; <= 32768, never carries between the halves)
        uint32_t RA = (ds << 2) | ((ds + S0 + S1) << 18);
        uint32_t RB = ((ds + S0) << 2) | ((ds + S0 + S1 + S2) << 18);
        uint32_t v;
#define COL_STEP(Aw, Bw) v = Aw; Aw = RA; RA += v << 2; v = Bw; Bw = RB; RB += v << 2;
        COL_STEP(r0.x, r0.y) COL_STEP(r0.z, r0.w) COL_STEP(r1.x, r1.y) COL_STEP(r1.z, r1.w)
        COL_STEP(r2.x, r2.y) COL_STEP(r2.z, r2.w) COL_STEP(r3.x, r3.y) COL_STEP(r3.z, r3.w)
#undef COL_STEP
        row[0] = r0; row[1] = r1; row[2] = r2; row[3] = r3;
    }
    __syncthreads();

    // payload loads are issued here: their latency hides behind the ranking, and their 32 registers
    // are not held through the count and the scans
    uint32_t val[HAS_VALUES ? I : 1];
    if (HAS_VALUES) {
        const uint32_t* src = vals_in + tile_base + t * I;
        if (FULL) {
#pragma unroll
            for (int v = 0; v < I / 8; v++) ld_v8(src + v * 8, val + v * 8);
        } else {
#pragma unroll
            for (int i = 0; i < I; i++) val[i] = (t * I + i) < valid ? src[i] : 0u;
        }
    }

    // ---- rank: slot = table[digit][my group] + (threads of my group before me: the scanned
    // counter) + (my earlier keys: occ); kept as the swizzled byte offset of the slot (x 2 for
    // key/value pairs), two per register
    uint32_t rk[I / 2];
    {
        // entry of (b, g) inside a row: word g * 2 + (b & 1), half b >> 1
        const uint32_t boff = g * 8 + (b & 1) * 4 + (b >> 1) * 2;
#pragma unroll
        for (int i = 0; i < I; i += 2) {
            uint32_t ph[2];
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const uint32_t a = cnt_addr<BYTE>(key[i + h], toff, shift, mask);
                const uint32_t q = (uint32_t)cnt[a] + ((occ[(i + h) >> 2] >> (8 * ((i + h) & 3))) & 0xffu);
                const uint32_t d = digit_of<BYTE>(key[i + h], shift, mask);
                const uint32_t bs4 = *reinterpret_cast<const uint16_t*>(base + d * BASE_ROW + boff);
                const uint32_t r4 = bs4 + (q << 2);
                ph[h] = r4 ^ ((r4 >> 5) & 0x7cu);        // 4 * (r ^ ((r >> 5) & 31))
            }
            rk[i >> 1] = (ph[0] | (ph[1] << 16)) << (HAS_VALUES ? 1 : 0);
        }
    }
    __syncthreads();    // the counters are dead: their 64 KB become the sorted tile

    // ---- scatter into the sorted tile
#pragma unroll
    for (int i = 0; i < I; i++) {
        const uint32_t off = (i & 1) ? (rk[i >> 1] >> 16) : (rk[i >> 1] & 0xffffu);
        if (HAS_VALUES) *reinterpret_cast<uint2*>(cnt + off) = make_uint2(key[i], val[i]);
        else *reinterpret_cast<uint32_t*>(cnt + off) = key[i];
    }

    // ---- decoupled look-back (thread = digit), 4 predecessors per round trip
    {
        constexpr int LB_WIDE = 4;
        uint32_t excl = 0;
        int64_t tt = (int64_t)tile - 1;
        bool done = tile == 0;
        while (!done) {
            uint32_t st[LB_WIDE];
#pragma unroll
            for (int k = 0; k < LB_WIDE; k++)
                st[k] = tt - k >= 0 ? ld_relaxed_u32(status + (size_t)(tt - k) * NDIG + t) : ST_INCLUSIVE;
#pragma unroll
            for (int k = 0; k < LB_WIDE; k++) {
                if (!done) {
                    while ((st[k] >> 30) == 0) st[k] = ld_relaxed_u32(status + (size_t)(tt - k) * NDIG + t);
                    excl += st[k] & ST_VALUE;
                    done = (st[k] >> 30) == 2;
                }
            }
            tt -= LB_WIDE;
        }
        if (tile > 0) st_relaxed_u32(status + (size_t)tile * NDIG + t, ST_INCLUSIVE | (excl + real));
        s_goff[t] = digit_base[t] + excl - ds;
    }
    __syncthreads();

    // ---- write out: consecutive threads -> consecutive slots -> runs of consecutive addresses
#pragma unroll
    for (int j = 0; j < I; j++) {
        const uint32_t pos = j * T + t;
        if (FULL || pos < valid) {
            const uint32_t sw = pos ^ ((j * (T / 32) + warp) & 31u);      // pos ^ ((pos >> 5) & 31)
            uint32_t k, v = 0;
            if (HAS_VALUES) {
                const uint2 kv = *reinterpret_cast<const uint2*>(cnt + sw * 8);
                k = kv.x;
                v = kv.y;
            } else {
                k = *reinterpret_cast<const uint32_t*>(cnt + sw * 4);
            }
            const uint32_t dst = s_goff[digit_of<BYTE>(k, shift, mask)] + pos;
#ifdef COL_DEBUG
            if (dst >= n) __trap();
#endif
            keys_out[dst] = k;
            if (HAS_VALUES) vals_out[dst] = v;
        }
    }
}

template <int BYTE, bool HAS_VALUES>
__global__ void __launch_bounds__(T, 2)
onesweep_col_kernel(const uint32_t* __restrict__ keys_in, uint32_t* __restrict__ keys_out,
                    const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out, uint32_t n,
                    const uint32_t* __restrict__ digit_base, uint32_t* __restrict__ status, unsigned* ticket,
                    uint32_t shift, uint32_t mask) {
    extern __shared__ __align__(16) unsigned char col_smem[];
    __shared__ uint32_t s_goff[NDIG];
    __shared__ uint32_t s_dstart[NDIG];
    __shared__ uint32_t s_warp[T / 32];
    __shared__ unsigned s_tile;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    {   // zero the counters
        uint4* z = reinterpret_cast<uint4*>(col_smem);
#pragma unroll
        for (int i = 0; i < CNT_BYTES / 16 / T; i++) z[i * T + threadIdx.x] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
    const unsigned tile = s_tile;
    if ((uint64_t)(tile + 1) * TILE <= n)
        col_tile<BYTE, true, HAS_VALUES>(keys_in, keys_out, vals_in, vals_out, n, digit_base, status, shift, mask, tile,
                                         col_smem, s_goff, s_dstart, s_warp);
    else
        col_tile<BYTE, false, HAS_VALUES>(keys_in, keys_out, vals_in, vals_out, n, digit_base, status, shift, mask, tile,
                                          col_smem, s_goff, s_dstart, s_warp);
}

template <int BYTE, bool HAS_VALUES>
int launch_variant(const uint32_t* kin, uint32_t* kout, const uint32_t* vin, uint32_t* vout, uint32_t n,
                   const uint32_t* digit_base, uint32_t* status, unsigned* ticket, uint32_t shift, uint32_t mask,
                   unsigned tiles, cudaStream_t s) {
    constexpr size_t smem = smem_bytes<HAS_VALUES>();
    COL_CUDA(cudaFuncSetAttribute(onesweep_col_kernel<BYTE, HAS_VALUES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    onesweep_col_kernel<BYTE, HAS_VALUES><<<tiles, T, smem, s>>>(kin, kout, vin, vout, n, digit_base, status, ticket, shift, mask);
    COL_CHECK_LAUNCH();
    return 0;
}

// one pass; the caller guarantees 32-byte aligned kin (and vin)
template <bool HAS_VALUES>
int launch(const uint32_t* kin, uint32_t* kout, const uint32_t* vin, uint32_t* vout, uint32_t n,
           const uint32_t* digit_base, uint32_t* status, unsigned* ticket, uint32_t shift, uint32_t mask,
           unsigned tiles, cudaStream_t s) {
    if (mask == 0xffu && (shift & 7u) == 0 && shift < 32) {
        switch (shift >> 3) {
            case 0: return launch_variant<0, HAS_VALUES>(kin, kout, vin, vout, n, digit_base, status, ticket, shift, mask, tiles, s);
            case 1: return launch_variant<1, HAS_VALUES>(kin, kout, vin, vout, n, digit_base, status, ticket, shift, mask, tiles, s);
            case 2: return launch_variant<2, HAS_VALUES>(kin, kout, vin, vout, n, digit_base, status, ticket, shift, mask, tiles, s);
            default: return launch_variant<3, HAS_VALUES>(kin, kout, vin, vout, n, digit_base, status, ticket, shift, mask, tiles, s);
        }
    }
    return launch_variant<-1, HAS_VALUES>(kin, kout, vin, vout, n, digit_base, status, ticket, shift, mask, tiles, s);
}

}  // namespace colsort
