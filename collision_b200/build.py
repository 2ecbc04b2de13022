"""Builds collision_b200/libcollision_b200.so from csrc/*.cu with nvcc for sm_100a.

The library is built IN-TREE (it travels to the GPU box with the repo snapshot)
and is the only compute path of the package: there is no fallback.

    python -m collision_b200.build [--force]
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
OBJ = HERE / "build"
LIB = HERE / "libcollision_b200.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    # IEEE arithmetic, identical to the reference's CPU evaluation: no FMA
    # contraction, correctly rounded division / sqrt, no flush-to-zero.
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (cand == "nvcc" or Path(cand).exists()):
            return cand
    return "nvcc"


def sources():
    return sorted(CSRC.glob("*.cu"))


def _stale(target, deps):
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps)


def build(force=False, verbose=False, debug=False):
    """Compile every .cu (in parallel) and link the shared library.  Returns its path.
    debug=True builds libcollision_b200_debug.so (-DCOL_DEBUG: traversal-stack depth traps;
    loaded instead of the release library when COL_B200_DEBUG=1 is set)."""
    global OBJ, LIB
    if debug:
        OBJ, LIB = HERE / "build_debug", HERE / "libcollision_b200_debug.so"
    headers = list(CSRC.glob("*.cuh")) + [HERE.parent / "include" / "collision_b200.h", Path(__file__)]
    OBJ.mkdir(exist_ok=True)
    nvcc = _nvcc()
    extra = os.environ.get("COL_NVCC_EXTRA", "").split()   # tuning aid (e.g. -DCOL_SORT_THREADS=256)
    if debug:
        extra.append("-DCOL_DEBUG")
    jobs = []
    for src in sources():
        obj = OBJ / (src.stem + ".o")
        if force or _stale(obj, [src] + headers):
            jobs.append([nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", str(src), "-o", str(obj)])

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: {}\n{}\n{}".format(" ".join(cmd), r.stdout, r.stderr))
        return r.stderr

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for log in ex.map(run, jobs):
            if verbose and log:
                print(log)
    objs = [OBJ / (s.stem + ".o") for s in sources()]
    if force or jobs or _stale(LIB, objs):
        run([nvcc, "-shared", "-o", str(LIB)] + [str(o) for o in objs] + ["-gencode", "arch=compute_100a,code=sm_100a"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, debug="--debug" in sys.argv))
