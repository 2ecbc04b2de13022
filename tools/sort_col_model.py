"""Bit-accurate numpy model of one tile of the column-counter onesweep pass
(collision_b200/csrc/sort_col.cuh): the same shared-memory layouts (byte counter per
(digit, thread) with four threads per word, 16-bit run-start table with 80-byte rows, XOR-swizzled
sorted tile) and the same packed arithmetic, so the index math -- incl. the corner where a group of
8 threads holds 256 keys of one digit -- can be checked without a GPU
(tests/test_sort_col_model.py).  Development aid; not imported by the package."""
import numpy as np

T, I, RADIX = 256, 32, 256
TILE = T * I
ROW = 80  # bytes per BASE row


def perm(b):
    return ((b & 1) << 1) | (b >> 1)


def swz(r):            # XOR swizzle of the sorted-tile index
    return r ^ ((r >> 5) & 31)


def tile_pass(keys, shift, mask, valid=None):
    """keys: TILE u32 (blocked: thread t holds keys[32t:32t+32]).  Returns (sorted tile, counts)."""
    valid = TILE if valid is None else valid
    keys = keys.copy()
    keys[valid:] = 0xFFFFFFFF
    dig = (keys >> shift) & mask
    cnt = np.zeros(65536, np.uint8)            # byte c[d][t] at d*256 + (t&63)*4 + (t>>6)
    t_idx = np.repeat(np.arange(T), I)
    xs = t_idx & 63
    toff = (xs ^ ((xs >> 5) << 2)) * 4 + (t_idx >> 6)      # halves of groups 4..7 swapped in memory
    # count phase (sequential per thread; numpy: accumulate)
    occ = np.zeros(TILE, np.int64)              # counter value before the key's own increment
    for idx in range(TILE):
        a = int(dig[idx]) * 256 + int(toff[idx])
        occ[idx] = cnt[a]
        cnt[a] += 1
    # scan A: unit u = d*8+g, 8 words
    words = cnt.view(np.uint32).copy()
    base = np.zeros(256 * ROW // 2, np.uint16)     # u16 entries, row stride 40 entries
    for u in range(2048):
        d, g = u >> 3, u & 7
        flip = (g >> 2) & 1
        w = np.concatenate([words[u * 8 + 4 * flip:u * 8 + 4 * flip + 4], words[u * 8 + 4 * (flip ^ 1):u * 8 + 4 * (flip ^ 1) + 4]]).astype(np.uint64)
        e = np.zeros(8, np.uint64)
        for j in range(1, 8):
            e[j] = (e[j - 1] + w[j - 1]) & 0xFFFFFFFF
        # no byte may overflow in e
        tlo = e[4]
        thi = (w[4] + w[5] + w[6] + w[7]) & 0xFFFFFFFF
        A = (tlo & 0x00ff00ff) + (thi & 0x00ff00ff)
        B = ((tlo >> 8) & 0x00ff00ff) + ((thi >> 8) & 0x00ff00ff)
        words[u * 8 + 4 * flip:u * 8 + 4 * flip + 4] = e[:4].astype(np.uint32)
        words[u * 8 + 4 * (flip ^ 1):u * 8 + 4 * (flip ^ 1) + 4] = e[4:].astype(np.uint32)
        ent = d * (ROW // 2) + g * 4
        base[ent + 0] = A & 0xffff
        base[ent + 1] = A >> 16
        base[ent + 2] = B & 0xffff
        base[ent + 3] = B >> 16
    cnt = words.view(np.uint8).copy()
    # scan B: thread d
    tot = np.zeros(256, np.int64)
    for d in range(256):
        row = base[d * (ROW // 2): d * (ROW // 2) + 32].astype(np.int64)
        tot[d] = row.sum()
    dstart = np.concatenate([[0], np.cumsum(tot)[:-1]])
    base4 = np.zeros_like(base)
    for d in range(256):
        row = base[d * (ROW // 2): d * (ROW // 2) + 32].astype(np.int64)
        S = [sum(row[g * 4 + perm(b)] for g in range(8)) for b in range(4)]
        sb = [dstart[d], dstart[d] + S[0], dstart[d] + S[0] + S[1], dstart[d] + S[0] + S[1] + S[2]]
        for b in range(4):
            run = sb[b]
            for g in range(8):
                base4[d * (ROW // 2) + g * 4 + perm(b)] = (4 * run) & 0xffff
                run += row[g * 4 + perm(b)]
    # rank phase
    out = np.zeros(TILE, np.uint32)
    phys = np.zeros(TILE, np.int64)
    for t in range(T):
        b, x = t >> 6, t & 63
        g = x >> 3
        for i in range(I):
            k = keys[t * I + i]
            d = int((k >> shift) & mask)
            a = d * 256 + (x ^ ((x >> 5) << 2)) * 4 + b
            q = int(cnt[a]) + int(occ[t * I + i])
            r4 = int(base4[d * (ROW // 2) + g * 4 + perm(b)]) + (q << 2)
            r = r4 >> 2
            phys[t * I + i] = swz(r)
    assert len(set(phys.tolist())) == TILE
    tile = np.zeros(TILE, np.uint32)
    tile[phys] = keys
    # write-out order
    pos = np.arange(TILE)
    sorted_tile = tile[swz(pos)]
    return sorted_tile, tot, dstart


def check(keys, shift, mask, valid=None):
    st, tot, ds = tile_pass(keys, shift, mask, valid)
    k = keys.copy()
    if valid is not None:
        k[valid:] = 0xFFFFFFFF
    order = np.argsort((k >> shift) & mask, kind='stable')
    assert np.array_equal(st, k[order]), "mismatch"
    assert np.array_equal(tot, np.bincount(((k >> shift) & mask).astype(np.int64), minlength=256))


def self_test():
    rs = np.random.RandomState(1)
    check(rs.randint(0, 1 << 32, TILE, dtype=np.uint64).astype(np.uint32), 8, 0xff)
    check(np.arange(TILE, dtype=np.uint32), 0, 0xff)
    check(np.arange(TILE, dtype=np.uint32), 8, 0xff)
    check(np.zeros(TILE, np.uint32), 0, 0xff)                      # everything in one bin: the overflow corners
    check(np.full(TILE, 0xffffffff, np.uint32), 24, 0xff)
    k = rs.randint(0, 4, TILE).astype(np.uint32)
    check(k, 0, 0xff)
    check(k, 0, 0xf)
    check(rs.randint(0, 1 << 30, TILE, dtype=np.int64).astype(np.uint32), 24, 0xff, valid=5000)   # partial tile
    check(np.zeros(TILE, np.uint32), 0, 0xff, valid=1)
    blk = np.repeat(rs.randint(0, 256, 32), 256).astype(np.uint32)   # 8 threads at a time share one digit
    check(blk, 0, 0xff)


if __name__ == "__main__":
    self_test()
    print("model ok")
