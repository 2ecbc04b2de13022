#!/bin/bash
# ONE compute-sanitizer tool per gpurun call (B200_PROFILING.md) over the small GPU tests that exercise the
# hand-rolled inter-block protocols (onesweep look-back, scan status words, refit hand-over, last-block fold).
#   bash tools/r2_sanitize.sh memcheck|racecheck|synccheck|initcheck TAG
TOOL=${1:-memcheck}; TAG=${2:-r2x}
O=gpurun_out
mkdir -p $O
SEL="random_collision or every_pass or prefix_sum or fused_build or test_bounds or configs0 or scan_golden or sorter_keys"
timeout 900 python -m pytest tests/test_collision_gpu.py tests/test_primitives_gpu.py -x -q -m gpu -k "$SEL and not 1048576 and not 70001 and not 4194309" > $O/${TAG}_${TOOL}_plain.log 2>&1 || { tail -5 $O/${TAG}_${TOOL}_plain.log; exit 1; }
timeout 2400 compute-sanitizer --tool $TOOL --error-exitcode 77 --log-file $O/${TAG}_sanitizer_${TOOL}.log \
   python -m pytest tests/test_collision_gpu.py tests/test_primitives_gpu.py -x -q -m gpu -k "$SEL and not 1048576 and not 70001 and not 4194309" > $O/${TAG}_${TOOL}_pytest.log 2>&1
echo "sanitizer rc=$?" >> $O/${TAG}_sanitizer_${TOOL}.log
tail -3 $O/${TAG}_${TOOL}_pytest.log; tail -8 $O/${TAG}_sanitizer_${TOOL}.log
