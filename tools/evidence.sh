#!/bin/bash
# Round evidence on one B200: tests, bench (both arms), stage timings, ncu launch list + full capture.
# usage (on the GPU box): bash tools/evidence.sh TAG    -> gpurun_out/TAG_*
TAG=${1:-rX}
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/${TAG}_tests.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_tests.log
timeout 600 python bench.py 2> $O/${TAG}_bench.err | tail -1 > $O/${TAG}_bench.json || exit 1
timeout 900 python bench.py --impl reference 2>> $O/${TAG}_bench.err | tail -1 > $O/${TAG}_bench_reference.json
timeout 600 python tools/stage_bench.py > $O/${TAG}_stage_bench.txt 2>&1
{ timeout 300 python tools/trv_bench.py; timeout 300 python tools/trv_bench.py --n 16777216 --clustered; timeout 300 python tools/trv_bench.py --n 134217728 --reps 2; } >> $O/${TAG}_stage_bench.txt 2>&1
# launch list (cold cache, serialised: shares only) -- the same command, after it exited 0 above
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches_bench.csv python bench.py --no-graph --steps 20 --warmup 5 > $O/${TAG}_ncu_launches.log 2>&1
