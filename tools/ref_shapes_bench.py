"""The reference's own benchmark shapes (tests/benchmarks/*.py upstream: pytest-benchmark over
PyOpenCL) through the mirrored Python API -- for comparison with anyone who has run upstream.
Same sizes, generators, group sizes and warm-up (10 rounds); device time per call from CUDA
events, median over the rounds (capped at 200).  SURVEY.md 8(f) f4.

    python tools/ref_shapes_bench.py            # prints one line per upstream benchmark case
"""
import sys
from functools import partial
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from collision_b200.bounds import Bounds                       # noqa: E402
from collision_b200.collision import Collider                  # noqa: E402
from collision_b200.offset import OffsetFinder                 # noqa: E402
from collision_b200.radix import RadixSorter                   # noqa: E402
from collision_b200.scan import PrefixScanner                  # noqa: E402

DEV = torch.device("cuda", 0)
WARMUP, MAX_ROUNDS = 10, 200


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).to(DEV)


def timed(name, fn, rounds, setup=None):
    cq = torch.cuda.current_stream()
    rounds = min(rounds, MAX_ROUNDS)
    ms = []
    for i in range(WARMUP + rounds):
        if setup:
            setup()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(cq)
        fn(cq)
        b.record(cq)
        b.synchronize()
        if i >= WARMUP:
            ms.append(a.elapsed_time(b))
    print("{:70s} median {:8.1f} us   min {:8.1f} us   ({} rounds)".format(name, 1e3 * float(np.median(ms)),
                                                                          1e3 * min(ms), rounds))


def bench_collide():   # tests/benchmarks/test_collide.py:25-58 (count-only call)
    for npoints in (307200, 307201):
        np.random.seed(4)
        coords = np.zeros((npoints, 4), np.float32)
        coords[:, :3] = np.random.uniform(-1.0, 1.0, (npoints, 3))
        radii = np.random.uniform(0.006, 0.06, npoints).astype(np.float32)
        cb, rb = dev(coords), dev(radii)
        n_buf = torch.zeros(1, dtype=torch.int32, device=DEV)
        col = Collider(DEV, npoints, 8, 128)
        timed("collide npoints={} rmax=0.06 (count only, as upstream)".format(npoints),
              lambda cq: col.get_collisions(cq, cb, rb, n_buf, None, 0), 10)
        total = int(n_buf.item()) & 0xFFFFFFFF
        pairs = torch.empty(max(total, 1) * 2, dtype=torch.int32, device=DEV)
        timed("collide npoints={} with the {} rows written".format(npoints, total),
              lambda cq: col.get_collisions(cq, cb, rb, n_buf, pairs, total), 10)


def bench_radix():     # tests/benchmarks/test_radix.py:51-115
    size = 307200
    gens = [("randint(0,1000)", partial(np.random.randint, 0, 1000)),
            ("randint(0,307200)", partial(np.random.randint, 0, 307200)), ("arange", np.arange)]
    for kdt in ("uint32", "uint64"):
        for gname, gen in gens:
            np.random.seed(4)
            keys = gen(size).astype(kdt)
            sorter = RadixSorter(DEV, size, 128, key_dtype=np.dtype(kdt))
            kin, kout = dev(keys), dev(np.zeros_like(keys))
            timed("radix_sort {} keys {} (keys only)".format(kdt, gname), lambda cq: sorter.sort(cq, kin, kout), 100)
    for vdt in (np.dtype("uint32"), np.dtype("float64"), np.dtype(("float32", 3)), np.dtype(("float32", 4))):
        np.random.seed(4)
        keys = np.random.randint(0, 307200, size).astype("uint32")
        lanes = 4 if vdt.shape == (3,) else int(np.prod(vdt.shape or (1,)))
        vals = np.random.random((size, lanes)).astype(vdt.base)
        sorter = RadixSorter(DEV, size, 128, key_dtype=np.dtype("uint32"), value_dtype=vdt)
        kin, kout, vin, vout = dev(keys), dev(np.zeros_like(keys)), dev(vals), dev(np.zeros_like(vals))
        timed("sort_values uint32 keys, values {}".format(vdt), lambda cq: sorter.sort(cq, kin, kout, vin, vout), 100)


def bench_scan():      # tests/benchmarks/test_scan.py:29-44
    for size, rounds in ((307200, 4000), (1536000, 800), (3072000, 400)):
        np.random.seed(4)
        src = dev(np.random.randint(0, 128, size).astype("uint32"))
        work = torch.empty_like(src)
        scanner = PrefixScanner(DEV, size, 128)
        timed("prefix_sum size={}".format(size), lambda cq: scanner.prefix_sum(cq, work), rounds,
              setup=lambda: work.copy_(src))


def bench_bounds():    # tests/benchmarks/test_bounds.py:18-36
    for size, rounds in ((1536000, 800), (3072000, 400)):
        np.random.seed(4)
        vals = dev(np.random.uniform(0.0, 1.0, (size, 4)).astype("float32"))
        out = torch.zeros(2 * 4 * 4, dtype=torch.uint8, device=DEV)
        red = Bounds(DEV, 64, 128)
        timed("bounds size={}".format(size), lambda cq: red.reduce(cq, size, vals, out), rounds)


def bench_offsets():   # tests/benchmarks/test_offset.py:24-44
    size = 2 ** 21
    for maxval in (2 * 10 ** 3, 2 * 10 ** 6):
        np.random.seed(4)
        values = dev(np.sort(np.random.randint(0, maxval, size, dtype="uint32")))
        offsets = torch.zeros(maxval * 4, dtype=torch.uint8, device=DEV)
        finder = OffsetFinder(DEV)
        timed("find_offsets size=2^21 maxval={}".format(maxval),
              lambda cq: finder.find_offsets(cq, values, size, offsets, maxval), 800)


if __name__ == "__main__":
    torch.cuda.set_device(DEV)
    print("# upstream benchmark shapes on", torch.cuda.get_device_name(DEV))
    bench_collide()
    bench_radix()
    bench_scan()
    bench_bounds()
    bench_offsets()
