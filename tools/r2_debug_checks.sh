#!/bin/bash
# compute-sanitizer is closed on this pool (B200_PROFILING.md: "a tool that answers that it is closed stays closed";
# the refusal is recorded in profiles/r2n_sanitizer_closed.txt).  In its place: the whole GPU suite against the
# -DCOL_DEBUG library (device traps on out-of-range sort destinations, impossible tree ranges / node indices,
# traversal-stack depth) with 4 KB canaries around every scratch buffer, checked after every test.
TAG=${1:-r2x}
O=gpurun_out; mkdir -p $O
COL_B200_DEBUG=1 timeout 2400 python -m pytest tests -x -q -m gpu > $O/${TAG}_debug_checks.log 2>&1; echo "pytest rc=$? (COL_B200_DEBUG=1: libcollision_b200_debug.so, scratch canaries)" >> $O/${TAG}_debug_checks.log
tail -4 $O/${TAG}_debug_checks.log
