"""Per-stage device timings (CUDA events on the launching stream, L2 flushed
between repetitions) with achieved GB/s against the algorithmic bytes of
SURVEY.md section 8(d).  Development aid; bench.py is the contract.

    python tools/stage_bench.py [--n 1048576] [--sort-n 67108864] [--reps 5]
"""
import argparse
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from collision_b200 import _lib  # noqa: E402
from collision_b200.collision import Collider, MORTON_BITS  # noqa: E402


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    return json.loads(p.read_text())["hbm_gbs"] if p.exists() else 6650.0


class Timer:
    def __init__(self, dev, reps):
        self.reps = reps
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2
        self.stream = torch.cuda.current_stream()

    def __call__(self, fn, setup=None):
        ts = []
        for _ in range(self.reps + 2):
            if setup:
                setup()
            self.flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            ts.append(a.elapsed_time(b))
        return float(np.median(ts[2:]))


def gen_uniform(n, dev, seed=4):
    g = torch.Generator(device=dev).manual_seed(seed)
    c = torch.zeros(n, 4, device=dev)
    c[:, :3] = torch.rand(n, 3, device=dev, generator=g)
    r = torch.full((n,), 0.5 * n ** (-1.0 / 3.0), device=dev)
    return c, r


def gen_clustered(n, dev, seed=4, k=1024, sigma=0.02):
    g = torch.Generator(device=dev).manual_seed(seed)
    centres = torch.rand(k, 3, device=dev, generator=g)
    which = torch.randint(0, k, (n,), device=dev, generator=g)
    c = torch.zeros(n, 4, device=dev)
    c[:, :3] = centres[which] + torch.randn(n, 3, device=dev, generator=g) * sigma
    r = torch.full((n,), 0.5 * n ** (-1.0 / 3.0), device=dev)
    return c, r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1 << 20)
    ap.add_argument("--sort-n", type=int, default=1 << 26)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--clustered", action="store_true")
    ap.add_argument("--skip-sort", action="store_true")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    L, P = _lib.lib(), _lib.ptr
    cq = torch.cuda.current_stream()
    S = cq.cuda_stream
    T = Timer(dev, args.reps)
    peak = peaks()
    rows = []

    def report(name, ms, nbytes):
        gbs = nbytes / ms / 1e6
        rows.append((name, ms, gbs, gbs / peak))
        print("{:34s} {:9.3f} ms {:9.1f} GB/s  {:5.1f}% of measured {:.0f} GB/s".format(name, ms, gbs, 100 * gbs / peak, peak), flush=True)

    n = args.n
    coords, radii = (gen_clustered if args.clustered else gen_uniform)(n, dev)
    col = Collider(dev, n, 8, 128)
    col._ensure_scratch()
    n_buf = torch.zeros(1, dtype=torch.int32, device=dev)
    code = 0
    rs = col.reducer._ensure_scratch()
    ss = col.sorter._ensure_scratch()

    report("bounds f32x4 n={}".format(n), T(lambda: L.col_bounds(P(coords), n, code, 4, P(col._bounds_buf), P(rs), S)), 16 * n)
    report("morton+ids", T(lambda: L.col_morton(P(col._codes_bufs[0]), P(col._ids_bufs[0]), P(coords), P(col._bounds_buf), code, n, n, S)), 24 * n)
    report("sort pairs (30 bit)", T(lambda: L.col_sort_pairs(P(col._codes_bufs[0]), P(col._codes_bufs[1]), P(col._ids_bufs[0]), P(col._ids_bufs[1]), n, 4, 4, MORTON_BITS, P(ss), S)), 68 * n)
    report("build tree", T(lambda: L.col_build_tree(P(col._codes_bufs[1]), P(col._ids_bufs[1]), P(col._nodes_buf), n, S)), 40 * n)

    def scene():
        L.col_bounds(P(coords), n, code, 4, P(col._bounds_buf), P(rs), S)
    report("leaf boxes + refit", T(lambda: L.col_refit(P(col._bounds_buf), P(col._flags_buf), P(col._nodes_buf), P(coords), P(radii), code, n, S)), 192 * n)
    report("traverse count", T(lambda: L.col_traverse_count(P(col._counts_buf), P(col._nodes_buf), P(col._bounds_buf), code, n, 0, n, S)), 36 * n)
    report("scan counts", T(lambda: L.col_exclusive_scan_u32(P(col._counts_buf), n, P(col._total_buf), P(col._scan_scratch), S),
                            setup=lambda: L.col_traverse_count(P(col._counts_buf), P(col._nodes_buf), P(col._bounds_buf), code, n, 0, n, S)), 8 * n)
    torch.cuda.synchronize()
    L.col_traverse_count(P(col._counts_buf), P(col._nodes_buf), P(col._bounds_buf), code, n, 0, n, S)
    L.col_exclusive_scan_u32(P(col._counts_buf), n, P(col._total_buf), P(col._scan_scratch), S)
    torch.cuda.synchronize()
    m = int(col._total_buf.item())
    print("pairs M = {} ({:.2f} per sphere)".format(m, m / n))
    pairs = torch.empty(max(m, 1) * 2, dtype=torch.int32, device=dev)
    scr_ = _lib.alloc_bytes(dev, L.col_traverse_scratch_bytes(n, m))
    report("traverse write", T(lambda: L.col_traverse_write(P(pairs), m, P(col._counts_buf), 0, P(col._nodes_buf), P(col._bounds_buf), code, n, 0, n, S)), 36 * n + 8 * m)
    report("traverse atomic (count only)", T(lambda: L.col_traverse_atomic(None, P(n_buf), 0, P(col._nodes_buf), P(col._bounds_buf), code, n, 0, n, S)), 36 * n)

    # ---- fast path (the Collider's default): 32-byte traversal records
    L.col_bounds(P(coords), n, code, 4, P(col._scene_buf), P(rs), S)
    report("build tree + wide meta", T(lambda: L.col_build_tree_wide(P(col._codes_bufs[1]), P(col._ids_bufs[1]), P(col._nodes_buf), P(col._wide_buf), P(col._flags_buf), n, S)), 60 * n)
    report("leaf boxes + refit (wide)", T(lambda: L.col_refit_wide(P(col._bounds_buf), P(col._wide_buf), P(col._flags_buf), P(col._nodes_buf), P(coords), P(radii), P(col._scene_buf), code, n, S),
                                          setup=lambda: L.col_build_tree_wide(P(col._codes_bufs[1]), P(col._ids_bufs[1]), P(col._nodes_buf), P(col._wide_buf), P(col._flags_buf), n, S)), 192 * n)
    report("tree + leaf boxes + refit, one climb", T(lambda: L.col_build_refit_wide(P(col._codes_bufs[1]), P(col._ids_bufs[1]), P(col._nodes_buf), P(col._wide_buf), P(col._flags_buf), P(col._bounds_buf), P(coords), P(radii), P(col._scene_buf), code, n, S)), 156 * n)
    report("traverse wide count", T(lambda: L.col_traverse_wide_count(P(col._total_buf), None, P(col._wide_buf), P(col._bounds_buf), P(col._scene_buf), code, n, 0, n, S)), 36 * n)
    report("traverse wide pairs (walk+scan+emit)", T(lambda: L.col_traverse_wide_pairs(P(pairs), m, 0, P(col._total_buf), P(col._wide_buf), P(col._bounds_buf), P(col._scene_buf), P(col._ids_bufs[1]), code, n, 0, n, P(scr_), S)), 36 * n + 8 * m)
    torch.cuda.synchronize()
    assert int(col._total_buf.item()) == m, (int(col._total_buf.item()), m)

    def full():
        col.get_collisions(cq, coords, radii, n_buf, pairs, m)
    report("FULL get_collisions (ordered)", T(full), 420 * n + 8 * m)
    ms = rows[-1][1]
    print("  -> {:.1f} M spheres/s, {:.1f} M pairs/s".format(n / ms / 1e3, m / ms / 1e3))

    if not args.skip_sort:
        del col, pairs
        torch.cuda.empty_cache()
        sn = args.sort_n
        g = torch.Generator(device=dev).manual_seed(4)
        scratch = _lib.alloc_bytes(dev, L.col_sort_scratch_bytes(sn, 4, 4))
        kout = torch.empty(sn, dtype=torch.int32, device=dev)
        vin = torch.arange(sn, dtype=torch.int32, device=dev)
        vout = torch.empty(sn, dtype=torch.int32, device=dev)
        for name, keys in (("uniform32", torch.randint(-2**31, 2**31 - 1, (sn,), dtype=torch.int32, device=dev, generator=g)),
                           ("uniform30", torch.randint(0, 2**30, (sn,), dtype=torch.int32, device=dev, generator=g)),
                           ("sorted", torch.arange(sn, dtype=torch.int32, device=dev))):
            report("sort keys  {} n={}".format(name, sn), T(lambda: L.col_sort_pairs(P(keys), P(kout), None, None, sn, 4, 0, 0, P(scratch), S)), 36 * sn)
            report("sort pairs {} n={}".format(name, sn), T(lambda: L.col_sort_pairs(P(keys), P(kout), P(vin), P(vout), sn, 4, 4, 0, P(scratch), S)), 68 * sn)
        vals = torch.randint(0, 128, (sn,), dtype=torch.int32, device=dev, generator=g)
        sscr = _lib.alloc_bytes(dev, L.col_scan_scratch_bytes(sn))
        report("scan u32 n={}".format(sn), T(lambda: L.col_exclusive_scan_u32(P(vals), sn, None, P(sscr), S)), 8 * sn)
        c16 = torch.rand(1 << 24, 4, device=dev, generator=g)
        out = torch.empty(8, device=dev)
        rscr = _lib.alloc_bytes(dev, L.col_reduce_scratch_bytes())
        report("bounds f32x4 n=16M", T(lambda: L.col_bounds(P(c16), 1 << 24, 0, 4, P(out), P(rscr), S)), 16 << 24)


if __name__ == "__main__":
    main()
