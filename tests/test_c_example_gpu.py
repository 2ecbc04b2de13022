"""examples/pipeline.c -- the hot path driven from C through the C ABI alone -- built with gcc,
loaded with ctypes and run on the device: same rows as the oracle."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np
import pytest

import golden_data as G

pytestmark = pytest.mark.gpu


class Buffers(C.Structure):
    _fields_ = [(name, C.c_void_p) for name in (
        "coords", "radii", "codes0", "codes1", "ids0", "ids1", "scene", "nodes", "bounds", "wide", "flags",
        "pairs", "total", "reduce_scratch", "sort_scratch", "traverse_scratch")]


def test_c_host_runs_the_pipeline(cuda_env, oracle, tmp_path):
    import torch
    from collision_b200 import _lib
    ROOT = Path(__file__).resolve().parent.parent
    dev, cq = cuda_env
    so = tmp_path / "pipeline.so"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-shared", "-fPIC", "-I", str(ROOT / "include"),
                    str(ROOT / "examples" / "pipeline.c"), "-o", str(so), "-L", str(_lib.lib_path().parent),
                    "-l:" + _lib.lib_path().name, "-Wl,-rpath," + str(_lib.lib_path().parent)], check=True)
    _lib.lib()                      # (libcollision_b200.so is then already mapped)
    ex = C.CDLL(str(so))
    n = 50000
    c, r = G.random_spheres(n)
    want = oracle.collide(c, r)
    cap = want["total"]
    sizes = [C.c_size_t(), C.c_size_t(), C.c_size_t()]
    ex.col_example_scratch(C.c_uint32(n), C.c_uint64(cap), *[C.byref(x) for x in sizes])
    sizes = [x.value for x in sizes]

    def raw(nbytes):
        return torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=dev)

    t = dict(coords=torch.from_numpy(np.ascontiguousarray(c, np.float32)).to(dev),
             radii=torch.from_numpy(np.ascontiguousarray(r, np.float32)).to(dev),
             codes0=raw(4 * n), codes1=raw(4 * n), ids0=raw(4 * n), ids1=raw(4 * n), scene=raw(32),
             nodes=raw(16 * (2 * n - 1)), bounds=raw(32 * (2 * n - 1)), wide=raw(32 * (n - 1)), flags=raw(8 * n),
             pairs=raw(8 * cap), total=raw(8), reduce_scratch=raw(sizes[0]), sort_scratch=raw(sizes[1]),
             traverse_scratch=raw(sizes[2]))
    b = Buffers(**{k: v.data_ptr() for k, v in t.items()})
    cq.wait_stream(torch.cuda.current_stream(dev))
    ex.col_example_frame.restype = C.c_int
    rc = ex.col_example_frame(C.byref(b), C.c_uint32(n), C.c_uint64(cap), C.c_void_p(cq.cuda_stream))
    assert rc == 0
    cq.synchronize()
    assert int(t["total"].view(torch.int64).item()) == want["total"]
    got = t["pairs"].cpu().numpy().view(np.uint32).reshape(-1, 2)
    np.testing.assert_equal(got, want["pairs"])
