#!/usr/bin/env python
"""bench.py -- the driver's benchmark contract for the LBVH sphere-collision path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one full Collider.get_collisions() over one synthetic scene:
scene bounds -> Morton codes -> stable radix sort -> Karras tree -> AABB refit ->
traversal -> deterministic pair compaction, writing the whole ordered pair list.

Workloads (BASELINE.json `configs`, float32 coordinates in 16-byte slots, seed 4):
  N = 1  : configs[3] -- 2^24 spheres in 1,024 Gaussian clusters (sigma 0.02), r = 0.5 * N^(-1/3):
           north_star's "full-pipeline 16M-sphere" case, ~243.7 M reported rows.  The 2^20
           uniform scene of configs[1] is timed beside it (`config_b`), and the standalone
           primitives at their target sizes (`stages_at_target`: configs[2]).
  N > 1  : configs[4] -- 2^27 uniform spheres, r = 0.5 * N^(-1/3), STRONG scaling: the scene
           arrives sharded by sphere index (rank r holds [n*r/N, n*(r+1)/N)), every rank
           traverses its contiguous Morton-sorted query range, per-rank row totals are
           all-gathered.  The weak-scaled configs[1] numbers (2^20 spheres per GPU) ride
           along as `weak_b`; `parity_check` compares every rank's rows with the 1-GPU list.

`value`  : spheres/s with the inputs already resident in HBM (device timed with CUDA events
           per step, L2 flushed between steps, max over ranks).
`e2e`    : the same metric through the public API with HOST (pinned) buffers, run as a
           stream of frames: every step copies its inputs host->device and its row count and
           rows device->host inside the timed region; the copies of neighbouring steps
           overlap the compute (double buffering).
`roofline`: the HBM-bound kernel furthest below the north_star target (the radix sort at
           2^26 pairs), timed live with CUDA events; the traversal -- the largest share of
           the step -- is latency-bound and reported under `traversal`.
`cpu_baseline` / `--impl reference`: the reference's algorithm on the host cores.  The
           reference itself needs PyOpenCL + an OpenCL CPU device (pocl): bench.py probes for
           them (`cpu_baseline.pocl_probe`) and, when they are absent, times the C/OpenMP
           port (oracle/collision_oracle.c: single walk + counter append, as upstream).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "spheres_per_sec_full_lbvh_pipeline"
UNIT = "spheres/s"
SEED = 4
NOMINAL_HBM_GBS = 8000.0     # north_star's "~8 TB/s"


# ------------------------------------------------------------------ scenes
def radius_for(n):
    return np.float32(0.5 * n ** (-1.0 / 3.0))


def uniform_scene(n, seed=SEED):
    """configs[1] / configs[4]: uniform in the unit cube, ~8 box-overlap neighbours each."""
    rs = np.random.RandomState(seed)
    c = np.zeros((n, 4), np.float32)
    c[:, :3] = rs.random_sample((n, 3))
    return c, np.full(n, radius_for(n), np.float32)


def clustered_scene(n, seed=SEED, k=1024, sigma=0.02, r_for=None):
    """configs[3]: k Gaussian clusters (sigma 0.02, not clipped); r from `r_for` spheres."""
    rs = np.random.RandomState(seed)
    centres = rs.random_sample((k, 3)).astype(np.float32)
    which = rs.randint(0, k, n)
    c = np.zeros((n, 4), np.float32)
    c[:, :3] = centres[which] + rs.standard_normal((n, 3)).astype(np.float32) * np.float32(sigma)
    return c, np.full(n, radius_for(r_for or n), np.float32)


WORKLOADS = {
    "B": dict(n=1 << 20, scene=uniform_scene,
              text="configs[1]: {n} uniform spheres in the unit cube, r=0.5*N^(-1/3) (~8 neighbours), "
                   "full pipeline incl. ordered pair list"),
    "D": dict(n=1 << 24, scene=clustered_scene,
              text="configs[3]: {n} spheres in 1024 Gaussian clusters (sigma 0.02), r=0.5*N^(-1/3) "
                   "(~14.5 rows per sphere), full pipeline incl. ordered pair list"),
    "E": dict(n=1 << 27, scene=None,
              text="configs[4]: {n} uniform spheres in the unit cube, r=0.5*N^(-1/3) (~8 neighbours), "
                   "full pipeline incl. ordered pair list, strong scaling"),
}


def workload_config(key, n, extra=None):
    cfg = {"workload": WORKLOADS[key]["text"].format(n=n), "n_spheres": n, "coord_dtype": "float32",
           "seed": SEED, "l2": "flushed between steps (256 MiB write)"}
    cfg.update(extra or {})
    return cfg


def measured_hbm_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line)

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------- the reference arm
def pocl_probe():
    """one probe for the reference's own runtime (PyOpenCL + a CPU
    OpenCL platform such as pocl).  Returns a short string recorded in the JSON line."""
    try:
        import pyopencl   # noqa: F401  (absent on these boxes)
    except Exception as e:   # ImportError, or a broken install
        return "import pyopencl failed: {}: {}".format(type(e).__name__, e)
    try:
        import pyopencl as cl
        plats = cl.get_platforms()
        cpus = [d.name for p in plats for d in p.get_devices() if d.type & cl.device_type.CPU]
        return "pyopencl {} present; CPU devices: {}".format(cl.VERSION_TEXT, cpus or "none")
    except Exception as e:
        return "pyopencl imported but no usable platform: {}: {}".format(type(e).__name__, e)


def reference_sample(key, world):
    """(n, scene(), description) of the bounded sample the CPU arm runs for workload `key`.
    D runs whole when `whole` (decided from the first timed step); E is always sampled: an
    eighth of the volume at the same density (2^24 spheres) -- the per-sphere work is the
    same, the tree three levels shallower."""
    if key == "B":
        n = WORKLOADS["B"]["n"] * world
        return n, lambda: uniform_scene(n), "the whole {}-sphere scene".format(n)
    if key == "D":
        n = WORKLOADS["D"]["n"]
        return n, lambda: clustered_scene(n), "the whole {}-sphere scene".format(n)
    n = 1 << 24
    return n, lambda: uniform_scene(n), ("{} uniform spheres at the density of configs[4] (an eighth of its "
                                         "volume; same rows per sphere, tree 3 levels shallower)".format(n))


def d_sample():
    """configs[3] cut down for slow hosts: 64 of the 1,024 clusters (2^20 spheres, same sigma, same r)."""
    n = 1 << 20
    return n, lambda: clustered_scene(n, k=64, r_for=WORKLOADS["D"]["n"]), \
        "{} spheres = 64 of the 1024 clusters of configs[3] (same sigma, same r, same rows per sphere)".format(n)


def time_cpu_oracle(key, world, steps, warmup, budget_s):
    """The oracle's full pipeline (orc_pipeline_f32_single: the reference's single walk with
    counter append, rows in arbitrary order as upstream) on all host cores; the sample shrinks
    if `steps + warmup` runs of it would not fit `budget_s`.  Returns a dict."""
    import oracle
    oracle.build()
    cores = oracle.set_num_threads()          # all cores, whatever OMP_NUM_THREADS says (torchrun exports 1)
    n, scene, what = reference_sample(key, world)
    coords, radii = scene()
    pipe = oracle.Pipeline(n)
    t0 = time.perf_counter()
    total, _ = pipe.run(coords, radii, 0, single_walk=True)
    first = time.perf_counter() - t0
    if key == "D" and first * 1.3 * (steps + warmup) > budget_s:
        n, scene, what = d_sample()
        coords, radii = scene()
        pipe = oracle.Pipeline(n)
        total, _ = pipe.run(coords, radii, 0, single_walk=True)
    rows = np.zeros((max(total, 1), 2), np.uint32)
    for _ in range(warmup):
        pipe.run(coords, radii, total, single_walk=True, pairs=rows)
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        pipe.run(coords, radii, total, single_walk=True, pairs=rows)
        ts.append(time.perf_counter() - t0)
    sec = float(np.mean(ts))
    return {"value": n / sec, "sec": sec, "n": n, "pairs": total, "cores": cores, "what": what}


def run_reference(args):
    """--impl reference: the reference's algorithm on the host cores, rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = max(1, args.gpus)
    key = args.workload or ("D" if world == 1 else "E")
    probe = pocl_probe()
    r = time_cpu_oracle(key, world, args.steps, args.warmup, budget_s=200.0)
    n_full = WORKLOADS[key]["n"] * (world if key == "B" else 1)
    sample = "every step = {}, full pipeline incl. pair list (single walk + counter append, as upstream)".format(r["what"])
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["sec"] * 1e3, "higher_is_better": True,
        "scaling": "weak" if world == 1 or key == "B" else "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": workload_config(key, n_full, {"l2": "n/a (CPU)"}),
        "pairs": r["pairs"], "pairs_per_sec": r["pairs"] / r["sec"],
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": sample,
                         "sample_spheres": r["n"], "pocl_probe": probe,
                         "build": "gcc -O3 -march=native -fopenmp -ffp-contract=off (oracle/Makefile)"},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------- our arm
STAGE_BYTES = {   # algorithmic bytes per launch (SURVEY.md 8(d) / DESIGN.md), n spheres, m pairs
    "bounds": lambda n, m: 16 * n,
    "morton": lambda n, m: 24 * n,
    "sort": lambda n, m: 68 * n,
    "tree": lambda n, m: 40 * n,
    "refit": lambda n, m: 192 * n,
    "tree+refit": lambda n, m: 156 * n,   # codes 4 + ids 4 + coords 16 + radii 4 read; nodes 32 + boxes 64 + records 32 written
    "traverse_count": lambda n, m: 36 * n,
    "scan": lambda n, m: 8 * n,
    "traverse_write": lambda n, m: 36 * n + 8 * m,
    "traverse_pairs": lambda n, m: 44 * n + 8 * m,   # traversal 36n + compaction 8n + 8m
}

# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full captures
# (profiles/*_ncu_full_summary.txt); filled in per kernel where a capture of the CURRENT kernel exists
NCU_TRAFFIC = json.loads((ROOT / "profiles" / "ncu_traffic.json").read_text()) \
    if (ROOT / "profiles" / "ncu_traffic.json").exists() else {}


def frac_pair(gbps, peak):
    return {"GBps": gbps, "frac": gbps / peak, "frac_of_nominal_8TBps": gbps / NOMINAL_HBM_GBS}


def stages_at_target(dev, cq, peak, collider_d=None, scene_d=None):
    """The standalone primitives at the sizes north_star's ">= 60 % of the HBM roofline" is
    judged on: sort (2^26 keys, three key distributions, keys-only and pairs), scan 2^26,
    bounds 2^24, and -- when the configs[3] collider is passed -- compaction (scan + emit)."""
    import torch
    from collision_b200 import _lib
    L, P, S = _lib.lib(), _lib.ptr, cq.cuda_stream
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = {}

    def timed(fn, reps=5, warm=2):
        ts = []
        for i in range(reps + warm):
            flush.fill_(0)
            cq.wait_stream(torch.cuda.current_stream())
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(cq)
            fn()
            b.record(cq)
            b.synchronize()
            if i >= warm:
                ts.append(a.elapsed_time(b))
        return float(np.median(ts))

    n = 1 << 26
    g = torch.Generator(device=dev).manual_seed(SEED)
    dists = {
        "uniform32": lambda: torch.randint(-(1 << 31), 1 << 31, (n,), dtype=torch.int64, device=dev, generator=g).to(torch.int32),
        "uniform30": lambda: torch.randint(0, 1 << 30, (n,), dtype=torch.int64, device=dev, generator=g).to(torch.int32),
        "sorted": lambda: torch.arange(n, dtype=torch.int32, device=dev),
    }
    vals = torch.arange(n, dtype=torch.int32, device=dev)
    kout, vout = torch.empty_like(vals), torch.empty_like(vals)
    scratch = _lib.alloc_bytes(dev, L.col_sort_scratch_bytes(n, 4, 4))
    for name, make in dists.items():
        keys = make()
        torch.cuda.synchronize()
        for value_bytes, label, per in ((4, "sort_pairs", 68), (0, "sort_keys", 36)):
            def run():
                _lib.check(L.col_sort_pairs(P(keys), P(kout), P(vals) if value_bytes else None,
                                            P(vout) if value_bytes else None, n, 4, value_bytes, 0, P(scratch), S),
                           "col_sort_pairs")
            ms = timed(run)
            e = {"n": n, "ms": ms, "algorithmic_bytes": per * n, "G_items_per_s": n / ms / 1e6}
            e.update(frac_pair(per * n / ms / 1e6, peak))
            out["{}_{}".format(label, name)] = e
        del keys
    del kout, vout, scratch
    data = (vals & 127).contiguous()
    sscr = _lib.alloc_bytes(dev, L.col_scan_scratch_bytes(n))
    ms = timed(lambda: _lib.check(L.col_exclusive_scan_u32(P(data), n, None, P(sscr), S), "col_exclusive_scan_u32"))
    e = {"n": n, "ms": ms, "algorithmic_bytes": 8 * n}
    e.update(frac_pair(8 * n / ms / 1e6, peak))
    out["scan_u32"] = e
    del data, sscr, vals
    nb = 1 << 24
    pts = torch.rand(nb, 4, device=dev, generator=g)
    mm = torch.empty(8, device=dev)
    rscr = _lib.alloc_bytes(dev, L.col_reduce_scratch_bytes())
    ms = timed(lambda: _lib.check(L.col_bounds(P(pts), nb, 0, 4, P(mm), P(rscr), S), "col_bounds"), reps=9)
    e = {"n": nb, "ms": ms, "algorithmic_bytes": 16 * nb}
    e.update(frac_pair(16 * nb / ms / 1e6, peak))
    out["bounds_f32x4"] = e
    del pts
    if collider_d is not None:
        # compaction at configs[3]: scan of the row counts + emit, between the marks of the traversal call
        coords_d, radii_d, n_buf, pairs_d, cap = scene_d
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        for ev in evs:
            ev.record(cq)       # (creates the handles)
        cq.synchronize()
        marks = (ctypes.c_void_p * 3)(*[ev.cuda_event for ev in evs[1:4]])
        collider_d._trv_marks = marks
        walk, confirm, scan, emit = [], [], [], []
        for i in range(5):
            flush.fill_(0)
            cq.wait_stream(torch.cuda.current_stream())
            collider_d.get_collisions(cq, coords_d, radii_d, n_buf, pairs_d, cap)
            evs[4].record(cq)
            evs[4].synchronize()
            if i >= 1:
                confirm.append(evs[1].elapsed_time(evs[2]))
                scan.append(evs[2].elapsed_time(evs[3]))
                emit.append(evs[3].elapsed_time(evs[4]))
        collider_d._trv_marks = None
        nd = collider_d.size
        ms = float(np.median(scan) + np.median(emit))
        nbytes = 8 * nd + 8 * cap
        e = {"n": nd, "rows": cap, "ms": ms, "scan_ms": float(np.median(scan)), "emit_ms": float(np.median(emit)),
             "confirm_ms": float(np.median(confirm)), "algorithmic_bytes": nbytes,
             "note": "scan of the per-query row counts + emit of the ordered rows (ids gathered); emit_ms includes the 4-byte counter copy"}
        e.update(frac_pair(nbytes / ms / 1e6, peak))
        out["compaction_configs3"] = e
    return out


def rows_checksum(rows_i64, base, torch):
    """Order-sensitive 64-bit checksum of a slice of the row list whose first row has global
    index `base`: sum over rows of mix(row) * (2 * index + 1), wrapping mod 2^64."""
    K = -7046029254386353131          # 0x9E3779B97F4A7C15 as int64
    acc, n, chunk = 0, rows_i64.numel(), 1 << 24
    for lo in range(0, n, chunk):
        v = rows_i64[lo:lo + chunk]
        h = v * K
        h = h ^ (h >> 31)
        idx = torch.arange(lo, lo + v.numel(), dtype=torch.int64, device=v.device) + base
        acc = (acc + int((h * (2 * idx + 1)).sum().item())) & 0xFFFFFFFFFFFFFFFF
    return acc


def time_steps(torch, cq, flush, step, steps, barrier):
    """K steps, one CUDA-event pair per step on the launching stream, L2 flushed in between."""
    evs = []
    barrier()
    for _ in range(steps):
        flush.fill_(0)
        cq.wait_stream(torch.cuda.current_stream())
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(cq)
        step()
        b.record(cq)
        evs.append((a, b))
        torch.cuda.current_stream().wait_stream(cq)
    barrier()
    return [a.elapsed_time(b) for a, b in evs]


def stage_times(torch, cq, flush, collider, step, reps, barrier):
    """Per-stage times from event marks between the stages (marks switch graph replay off)."""
    collider.stage_events = []
    for _ in range(reps):
        flush.fill_(0)
        cq.wait_stream(torch.cuda.current_stream())
        step()
        torch.cuda.current_stream().wait_stream(cq)
    barrier()
    marks = collider.stage_events
    collider.stage_events = None
    ms = {}
    for (n0, e0), (n1, e1) in zip(marks, marks[1:]):
        if n1 != "begin":
            ms.setdefault(n1, []).append(e0.elapsed_time(e1))
    return {k: float(np.mean(v[1:] if len(v) > 1 else v)) for k, v in ms.items()}


def run_single(args):
    import torch
    from collision_b200 import _lib
    from collision_b200.collision import Collider

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: collision_b200 has no CPU fallback")
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    _lib.lib()   # fail loudly if the sm_100a library is missing
    key = args.workload or "D"
    n = WORKLOADS[key]["n"] if args.n_log2 is None else 1 << args.n_log2
    peak, peak_src = measured_hbm_peak()
    cq = torch.cuda.Stream(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    warm = max(args.warmup, 3)

    def barrier():
        torch.cuda.synchronize()

    def prepare(scene, n_s, graph=True):
        coords_h, radii_h = scene
        coords_pin = torch.from_numpy(coords_h).pin_memory()
        radii_pin = torch.from_numpy(radii_h).pin_memory()
        coords_d, radii_d = coords_pin.to(dev), radii_pin.to(dev)
        col = Collider(dev, n_s, 8, 128, graph=graph and not args.no_graph)
        n_buf = torch.zeros(1, dtype=torch.int32, device=dev)
        # size the pair buffer from a count-only call (the reference's own protocol)
        cq.wait_stream(torch.cuda.current_stream())
        col.get_collisions(cq, coords_d, radii_d, n_buf, None, 0).synchronize()
        total = int(n_buf.item()) & 0xFFFFFFFF
        pairs_d = torch.empty(max(total, 1) * 2, dtype=torch.int32, device=dev)
        return dict(col=col, coords_pin=coords_pin, radii_pin=radii_pin, coords_d=coords_d, radii_d=radii_d,
                    n_buf=n_buf, pairs_d=pairs_d, total=total, n=n_s)

    scene_fn = WORKLOADS[key]["scene"] or uniform_scene
    W = prepare(scene_fn(n), n)
    col, total = W["col"], W["total"]

    def step():
        return col.get_collisions(cq, W["coords_d"], W["radii_d"], W["n_buf"], W["pairs_d"], total)

    # clocks are sampled from the warm-up through the timed region
    sampler = ClockSampler(0)
    sampler.start()
    t_warm = time.perf_counter()
    for _ in range(warm):
        flush.fill_(0)
        step()
        torch.cuda.synchronize()
    per_step = (time.perf_counter() - t_warm) / warm
    for _ in range(min(2000, int(max(0.0, 0.5 - (time.perf_counter() - t_warm)) / max(per_step, 1e-5)))):
        flush.fill_(0)
        step()
    torch.cuda.synchronize()

    launches0 = col.kernel_launches
    step_ms = time_steps(torch, cq, flush, step, args.steps, barrier)
    clocks = sampler.stop()
    launches = col.kernel_launches - launches0
    ms_per_step = float(sum(step_ms)) / args.steps
    st = stage_times(torch, cq, flush, col, step, max(4, args.steps // 4), barrier)

    # ---- end to end through the public API with host buffers -----------------------------
    # (1) one step at a time: H2D of the inputs, get_collisions, D2H of the count, then of
    #     exactly that many rows; nothing overlaps -> the latency of one query from host memory
    pairs_pin = torch.empty(max(total, 1) * 2, dtype=torch.int32).pin_memory()
    n_pin = torch.zeros(1, dtype=torch.int32).pin_memory()
    lat_s = []
    lat_reps = min(args.steps, 5)
    for i in range(2 + lat_reps):
        flush.fill_(0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        with torch.cuda.stream(cq):
            W["coords_d"].copy_(W["coords_pin"], non_blocking=True)
            W["radii_d"].copy_(W["radii_pin"], non_blocking=True)
            step()
            n_pin.copy_(W["n_buf"], non_blocking=True)
        cq.synchronize()
        got = int(n_pin.item()) & 0xFFFFFFFF
        with torch.cuda.stream(cq):
            pairs_pin[:got * 2].copy_(W["pairs_d"][:got * 2], non_blocking=True)
        cq.synchronize()
        if i >= 2:
            lat_s.append(time.perf_counter() - t0)
    e2e_latency = float(np.mean(lat_s))

    # (2) the same work as a stream of frames, the way a simulation loop calls it: every step
    #     still copies its own inputs from pinned host memory and reads its own result (count +
    #     the whole row buffer) back, but step i+1's upload and step i-1's download run on
    #     their own streams while step i computes (double-buffered, ordered with the events
    #     get_collisions takes / returns).  K steps, wall clock / K.
    up, down = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    coords_db = [W["coords_d"], torch.empty_like(W["coords_d"])]
    radii_db = [W["radii_d"], torch.empty_like(W["radii_d"])]
    pairs_db = [W["pairs_d"], torch.empty_like(W["pairs_d"])]
    nbuf_db = [W["n_buf"], torch.zeros_like(W["n_buf"])]
    pairs_pin_db = [pairs_pin, torch.empty_like(pairs_pin).pin_memory()]
    n_pin_db = [n_pin, torch.zeros_like(n_pin).pin_memory()]
    computed, downloaded = [None, None], [None, None]

    def frames(k):
        for i in range(k):
            b = i & 1
            with torch.cuda.stream(up):
                if computed[b] is not None:
                    up.wait_event(computed[b])          # inputs of step i-2 no longer needed
                coords_db[b].copy_(W["coords_pin"], non_blocking=True)
                radii_db[b].copy_(W["radii_pin"], non_blocking=True)
                uploaded = torch.cuda.Event()
                uploaded.record(up)
            if downloaded[b] is not None:
                downloaded[b].synchronize()             # host has consumed step i-2's rows (back-pressure)
            ev = col.get_collisions(cq, coords_db[b], radii_db[b], nbuf_db[b], pairs_db[b], total, wait_for=[uploaded])
            computed[b] = ev
            with torch.cuda.stream(down):
                down.wait_event(ev)
                n_pin_db[b].copy_(nbuf_db[b], non_blocking=True)
                pairs_pin_db[b].copy_(pairs_db[b], non_blocking=True)
                d = torch.cuda.Event()
                d.record(down)
                downloaded[b] = d
        for d in downloaded:
            if d is not None:
                d.synchronize()

    frames(3)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    frames(args.steps)
    torch.cuda.synchronize()
    e2e_sec = (time.perf_counter() - t0) / args.steps
    assert (int(n_pin_db[0].item()) & 0xFFFFFFFF) == (total & 0xFFFFFFFF)
    del coords_db, radii_db, pairs_db, pairs_pin_db, pairs_pin

    # ---- the primitives at their target sizes + compaction at this workload ---------------
    sat = None
    if not args.no_stages:
        sat = stages_at_target(dev, cq, peak, col if key == "D" else None,
                               (W["coords_d"], W["radii_d"], W["n_buf"], W["pairs_d"], total))

    # ---- configs[1] beside it (continuity with round 1's headline) ------------------------
    config_b = None
    if key != "B" and not args.no_stages:
        nb = WORKLOADS["B"]["n"]
        B = prepare(uniform_scene(nb), nb)

        def step_b():
            return B["col"].get_collisions(cq, B["coords_d"], B["radii_d"], B["n_buf"], B["pairs_d"], B["total"])
        for _ in range(5):
            step_b()
        torch.cuda.synchronize()
        b_ms = time_steps(torch, cq, flush, step_b, max(args.steps, 20), barrier)
        b_st = stage_times(torch, cq, flush, B["col"], step_b, 8, barrier)
        bm = float(np.mean(b_ms))
        config_b = {"workload": WORKLOADS["B"]["text"].format(n=nb), "value": nb / (bm * 1e-3), "unit": UNIT,
                    "ms_per_step": bm, "pairs": B["total"], "pairs_per_sec": B["total"] / (bm * 1e-3),
                    "launch": "CUDA graph replay" if not args.no_graph else "kernel by kernel",
                    "stages_ms": b_st}
        del B

    # ---- CPU baseline: a bounded sample of the same workload on the host cores -------------
    cpu = None
    if not args.no_cpu:
        r = time_cpu_oracle(key, 1, 3, 1, budget_s=40.0)
        if r["n"] == n:
            assert r["pairs"] == total, ("GPU and oracle disagree on the pair count", total, r["pairs"])
        cpu = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port",
               "sample": "3 runs of {} (full pipeline incl. pair list; single walk + counter append as upstream), "
                         "oracle/collision_oracle.c with OpenMP".format(r["what"]),
               "sample_spheres": r["n"], "ms_per_run": r["sec"] * 1e3, "pocl_probe": pocl_probe()}

    # ---- roofline: the HBM-bound kernel furthest below the north_star target ---------------
    roofline = None
    if sat:
        hb = {k: v for k, v in sat.items() if k.startswith(("sort_", "scan_", "bounds_", "compaction"))}
        worst = min(hb, key=lambda k: hb[k]["frac"])
        e = hb[worst]
        roofline = {"kernel": worst, "bound": "hbm", "achieved": e["GBps"], "peak": peak, "unit": "GB/s",
                    "frac": e["frac"], "traffic": NCU_TRAFFIC.get(worst), "peak_source": peak_src,
                    "algorithmic_bytes": e["algorithmic_bytes"], "ms": e["ms"],
                    "note": "furthest below north_star's >= 60 % among the HBM-bound stages (sort, scan, reduce, "
                            "compaction; all in stages_at_target); per-launch = one whole call of the primitive"}
    trav_key = "traverse_pairs"
    traversal = None
    if trav_key in st:
        nbytes = STAGE_BYTES[trav_key](n, total)
        traversal = {"kernels": "candidates_kernel + confirm_kernel + scan_kernel + emit_kernel (one C-ABI call)",
                     "bound": "latency (dependent L1/L2 loads + lane divergence; see DESIGN.md section 4)",
                     "ms": st[trav_key], "share_of_step": st[trav_key] / ms_per_step,
                     "algorithmic_bytes": nbytes, "GBps": nbytes / st[trav_key] / 1e6,
                     "frac_of_hbm_peak": nbytes / st[trav_key] / 1e6 / peak,
                     "traffic": NCU_TRAFFIC.get("traverse_pairs_" + key)}
    if roofline is None and traversal is not None:
        roofline = {"kernel": trav_key, "bound": "latency", "achieved": traversal["GBps"], "peak": peak,
                    "unit": "GB/s", "frac": traversal["frac_of_hbm_peak"], "traffic": traversal["traffic"],
                    "peak_source": peak_src, "algorithmic_bytes": traversal["algorithmic_bytes"], "ms": traversal["ms"]}
    stages = {k: {"ms": v, "GBps": STAGE_BYTES.get(k, lambda a, b: 0)(n, total) / v / 1e6 if v > 0 else None}
              for k, v in st.items()}
    line = {
        "metric": METRIC, "value": n / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": 1,
        "steps": args.steps, "warmup": warm, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(key, n, {"launch": "CUDA graph replay" if not args.no_graph else "kernel by kernel",
                                           "parallelism": "single GPU"}),
        "pairs": total, "pairs_per_sec": total / (ms_per_step * 1e-3),
        "e2e": {"value": n / e2e_sec, "unit": UNIT,
                "h2d_bytes_per_step": W["coords_pin"].numel() * 4 + W["radii_pin"].numel() * 4,
                "d2h_bytes_per_step": 4 + total * 8, "ms_per_step": e2e_sec * 1e3,
                "how": "stream of frames: upload / compute / download of consecutive steps overlap "
                       "(double-buffered), every step copies its inputs H2D and its count + rows D2H",
                "single_step_latency_ms": e2e_latency * 1e3},
        "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "traversal": traversal,
        "stages": stages, "stages_at_target": sat, "config_b": config_b, "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)


_T0 = time.perf_counter()


def crumb(what):
    """Progress marks on stderr (rank 0): where a long multi-GPU run spends its wall clock."""
    if int(os.environ.get("RANK", "0")) == 0:
        print("[bench +{:6.1f}s] {}".format(time.perf_counter() - _T0, what), file=sys.stderr, flush=True)


def run_multi(args):
    import datetime

    import torch
    import torch.distributed as dist
    from collision_b200 import _lib
    from collision_b200.collision import Collider
    from collision_b200.sharded import ShardedCollider, shard_range

    world = int(os.environ["WORLD_SIZE"])
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: collision_b200 has no CPU fallback")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=300))
    _lib.lib()
    key = args.workload or "E"
    peak, peak_src = measured_hbm_peak()
    cq = torch.cuda.Stream(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    warm = max(args.warmup, 3)
    mode = args.multi_gpu_mode

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def device_scene(n, seed):
        """The whole scene, generated identically on every rank's GPU."""
        if key == "D" and seed == SEED:
            c, r = clustered_scene(n)
            return torch.from_numpy(c).to(dev), torch.from_numpy(r).to(dev)
        g = torch.Generator(device=dev).manual_seed(seed)
        c = torch.zeros(n, 4, device=dev)
        c[:, :3] = torch.rand(n, 3, device=dev, generator=g)
        return c, torch.full((n,), float(radius_for(n)), device=dev)

    def measure(n, seed, steps, with_e2e, with_parity):
        crumb("scene n={}".format(n))
        full_c, full_r = device_scene(n, seed)
        b, e = shard_range(n, world, rank)
        mine_pin_c = full_c[b:e].cpu().pin_memory()
        mine_pin_r = full_r[b:e].cpu().pin_memory()
        # what the rank holds when a step starts: its own slice (mode gather_inputs / sharded_build),
        # rank 0 everything (the broadcast modes); the rest is poison until the step fills it
        if mode in ("gather_inputs", "sharded_build"):
            coords_d = torch.full_like(full_c, float("nan"))
            radii_d = torch.full_like(full_r, float("nan"))
            coords_d[b:e] = full_c[b:e]
            radii_d[b:e] = full_r[b:e]
        else:
            coords_d, radii_d = (full_c.clone(), full_r.clone()) if rank == 0 else \
                (torch.full_like(full_c, float("nan")), torch.full_like(full_r, float("nan")))
        crumb("collider + count-only call")
        col = ShardedCollider(dev, n, 8, 128, mode=mode, graph=not args.no_graph, sharded_sort=args.shard_sort)
        n_buf = torch.zeros(1, dtype=torch.int32, device=dev)
        cq.wait_stream(torch.cuda.current_stream())
        col.get_collisions(cq, coords_d, radii_d, n_buf, None, 0).synchronize()
        total = col.total_pairs
        cap = col.local_pairs
        pairs_d = torch.empty(max(cap, 1) * 2, dtype=torch.int32, device=dev)

        def step():
            return col.get_collisions(cq, coords_d, radii_d, n_buf, pairs_d, cap, sync_totals=False)
        for _ in range(warm):
            flush.fill_(0)
            step()
            torch.cuda.synchronize()
        crumb("timed steps")
        launches0 = col.kernel_launches
        ms = time_steps(torch, cq, flush, step, steps, barrier)
        launches = col.kernel_launches - launches0
        st = stage_times(torch, cq, flush, col, step, 4, barrier)
        res = dict(total_ms=float(sum(ms)), total=total, cap=cap, launches=launches, stages=st, n=n)
        # one more call WITH the host bookkeeping: local_pairs / pair_base of the timed configuration
        col.get_collisions(cq, coords_d, radii_d, n_buf, pairs_d, cap).synchronize()
        res["sort_n"] = getattr(col, "sort_count", n)      # pairs this rank's sort moved (all of them unless the sort is dealt out)
        res["sort_mode"] = col.sort_mode
        assert col.local_pairs == cap and col.total_pairs == total
        assert (int(n_buf.item()) & 0xFFFFFFFF) == (total & 0xFFFFFFFF)

        if with_e2e:
            crumb("e2e")
            pin_rows = torch.empty(max(cap, 1) * 2, dtype=torch.int32).pin_memory()
            n_pin = torch.zeros(1, dtype=torch.int32).pin_memory()

            sharded_in = mode in ("gather_inputs", "sharded_build")
            if not sharded_in and rank == 0:
                all_pin_c, all_pin_r = full_c.cpu().pin_memory(), full_r.cpu().pin_memory()

            def frame():
                with torch.cuda.stream(cq):
                    if sharded_in:          # k host links feed the scene
                        coords_d[b:e].copy_(mine_pin_c, non_blocking=True)
                        radii_d[b:e].copy_(mine_pin_r, non_blocking=True)
                    elif rank == 0:         # the broadcast modes: rank 0 uploads everything
                        coords_d.copy_(all_pin_c, non_blocking=True)
                        radii_d.copy_(all_pin_r, non_blocking=True)
                    step()
                    n_pin.copy_(n_buf, non_blocking=True)
                    pin_rows.copy_(pairs_d, non_blocking=True)
            for _ in range(2):
                frame()
            torch.cuda.synchronize()
            barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                frame()
            torch.cuda.synchronize()
            res["e2e_sec"] = (time.perf_counter() - t0) / steps
            res["h2d"] = (mine_pin_c.numel() + mine_pin_r.numel()) * 4
            res["d2h"] = 4 + cap * 8
            del pin_rows

        if with_parity:
            crumb("parity check")
            # every rank's rows against the matching slice of the 1-GPU list (rank 0 runs the
            # single-GPU Collider on the same scene, outside every timed region)
            mine = torch.tensor([rows_checksum(pairs_d[:cap * 2].view(torch.int64), col.pair_base, torch) - (1 << 63),
                                 col.pair_base, cap], dtype=torch.int64, device=dev)
            every = torch.empty(world * 3, dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(every, mine)
            ok, n1 = None, None
            if rank == 0:
                one = Collider(dev, n, 8, 128)
                nb1 = torch.zeros(1, dtype=torch.int32, device=dev)
                rows1 = torch.empty(max(total, 1) * 2, dtype=torch.int32, device=dev)
                one.get_collisions(cq, full_c, full_r, nb1, rows1, total).synchronize()
                t1 = [0.0]
                if args.time_n1:
                    def step1():
                        return one.get_collisions(cq, full_c, full_r, nb1, rows1, total)
                    t1 = time_steps(torch, cq, flush, step1, 3, torch.cuda.synchronize)
                total1 = int(one._total_buf.item())
                ok = total1 == total
                v = rows1.view(torch.int64)
                for r in range(world):
                    cs, base, cnt = [int(x) for x in every[3 * r:3 * r + 3].tolist()]
                    ok = ok and rows_checksum(v[base:base + cnt], base, torch) - (1 << 63) == cs
                n1 = float(np.mean(t1))
                del one, rows1
            flag = torch.tensor([1 if ok else 0, 0], dtype=torch.int64, device=dev)
            dist.broadcast(flag, src=0)
            res["parity"] = bool(flag[0].item())
            res["n1_ms"] = n1
        col.close()
        del col, pairs_d, coords_d, radii_d, full_c, full_r
        torch.cuda.empty_cache()
        return res

    crumb("process group up")
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    n = WORKLOADS[key]["n"] if args.n_log2 is None else 1 << args.n_log2
    if key == "B":
        n *= world
    R = measure(n, 27 if key == "E" else SEED, args.steps, True, True)
    clocks = sampler.stop() if rank == 0 else None
    weak = None
    if key == "E" and not args.no_stages:
        weak = measure(WORKLOADS["B"]["n"] * world, SEED + 1, max(args.steps, 20), False, True)

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])
    total_ms = allmax(R["total_ms"])
    e2e_sec = allmax(R["e2e_sec"])
    h2d = n * 20        # every sphere's 16-byte slot + radius crosses a host link once per step
    d2h = R["total"] * 8 + 4 * world
    stages = {k: allmax(v) for k, v in sorted(R["stages"].items())}
    if weak is not None:
        weak_ms = allmax(weak["total_ms"]) / max(args.steps, 20)
        weak_stages = {k: allmax(v) for k, v in sorted(weak["stages"].items())}
    ms_per_step = total_ms / args.steps
    if rank == 0:
        trav = stages.get("traverse_pairs")
        traversal = None
        if trav:
            nbytes = STAGE_BYTES["traverse_pairs"](n // world, R["cap"])
            traversal = {"bound": "latency (dependent L1/L2 loads + lane divergence)", "ms": trav,
                         "share_of_step": trav / ms_per_step, "algorithmic_bytes": nbytes,
                         "GBps": nbytes / trav / 1e6, "frac_of_hbm_peak": nbytes / trav / 1e6 / peak,
                         "note": "this rank's 1/{} of the queries".format(world)}
        # the largest HBM-bound stage (the traversal and the build are latency-bound: `traversal` above)
        dom = max((k for k in stages if k in STAGE_BYTES and not k.startswith(("traverse", "tree"))),
                  key=lambda k: stages[k], default=None)
        roofline = None
        if dom:
            q = R["sort_n"] if dom == "sort" else n
            m = R["total"]
            nbytes = STAGE_BYTES[dom](q, m)
            roofline = {"kernel": dom, "bound": "hbm",
                        "achieved": nbytes / stages[dom] / 1e6, "peak": peak, "unit": "GB/s",
                        "frac": nbytes / stages[dom] / 1e6 / peak, "traffic": None, "peak_source": peak_src,
                        "algorithmic_bytes": nbytes, "ms": stages[dom], "share_of_step": stages[dom] / ms_per_step}
        line = {
            "metric": METRIC, "value": n / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": warm, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak" if key == "B" else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(key, n, {
                "launch": "kernel by kernel" if args.no_graph else "CUDA graph replay where the mode allows it",
                "parallelism": "{} ({} ranks; sort: {}): traversal sharded over contiguous Morton-sorted query ranges, "
                               "NCCL all_gather of the row totals".format(mode, world, R["sort_mode"]),
                "sorted_pairs_on_rank0": R["sort_n"]}),
            "pairs": R["total"], "pairs_per_sec": R["total"] / (ms_per_step * 1e-3),
            "e2e": {"value": n / e2e_sec, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_sec * 1e3,
                    "how": "every rank uploads its slice of the scene from pinned host memory and downloads its "
                           "count + its shard of the rows, every step"},
            "gpu_launches": R["launches"], "clocks": clocks, "roofline": roofline, "traversal": traversal,
            "stages": {k: {"ms": v} for k, v in stages.items()},
            "parity_check": {"rows_equal_single_gpu": R["parity"], "mode": mode,
                             "how": "order-sensitive 64-bit checksum of every rank's shard against the matching "
                                    "slice of the 1-GPU Collider's list on the same scene + equal totals"},
            "n1_same_workload": None if not R.get("n1_ms") else
            {"ms_per_step": R["n1_ms"], "value": n / (R["n1_ms"] * 1e-3),
             "speedup": R["n1_ms"] / ms_per_step, "efficiency": R["n1_ms"] / ms_per_step / world},
            "weak_b": None if weak is None else {
                "workload": WORKLOADS["B"]["text"].format(n=weak["n"]), "n_spheres": weak["n"],
                "ms_per_step": weak_ms, "value": weak["n"] / (weak_ms * 1e-3), "pairs": weak["total"],
                "stages_ms": weak_stages, "parity_check": {"rows_equal_single_gpu": weak["parity"], "mode": mode}},
        }
        print(json.dumps(line), flush=True)
    crumb("done")
    ok = R["parity"] and (weak is None or weak["parity"])
    dist.barrier()
    dist.destroy_process_group()
    if not ok:
        raise SystemExit("parity_check failed: the sharded rows differ from the single-GPU list")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default=None,
                    help="override the workload (default: D = configs[3] on one GPU, E = configs[4] on several)")
    ap.add_argument("--n-log2", type=int, default=None, help="override the scene size (development / CI)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--shard-sort", default=None, choices=["exchange", "local", "replicated"],
                    help="mode sharded_build: how the sort is dealt out (see ShardedCollider)")
    ap.add_argument("--no-stages", action="store_true", help="skip stages_at_target / config_b / weak_b (profiling runs)")
    ap.add_argument("--no-graph", action="store_true", help="enqueue kernel by kernel instead of replaying a CUDA graph")
    ap.add_argument("--time-n1", action="store_true", default=True,
                    help="N>1: also time the single-GPU Collider on the same scene (rank 0, outside the timed region)")
    ap.add_argument("--multi-gpu-mode", choices=["broadcast_tree", "replicate_build", "gather_inputs", "sharded_build"],
                    default="sharded_build",
                    help="N>1: sharded_build (default: build dealt out, other ranks' subtrees fetched on demand over "
                         "NVLink), gather_inputs / replicate_build (build replicated), broadcast_tree (north_star's "
                         "literal form: rank 0 builds, the tree is broadcast)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif int(os.environ.get("WORLD_SIZE", "1")) > 1:
        run_multi(args)
    else:
        run_single(args)


if __name__ == "__main__":
    main()
