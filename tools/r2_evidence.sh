#!/bin/bash
# Round-2 evidence on ONE B200: GPU tests, bench (both arms), then the ncu launch list of the same
# bench command (cold cache, serialised: compare SHARES).   usage: bash tools/r2_evidence.sh TAG [quick]
TAG=${1:-r2x}
O=gpurun_out
mkdir -p $O
if [ "$2" != "quick" ]; then
  timeout 2400 python -m pytest tests -x -q -m gpu > $O/${TAG}_tests.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_tests.log
fi
timeout 900 python bench.py 2> $O/${TAG}_bench.err | tail -1 > $O/${TAG}_bench.json || { echo "bench failed"; tail -20 $O/${TAG}_bench.err; exit 1; }
timeout 900 python bench.py --impl reference 2>> $O/${TAG}_bench.err | tail -1 > $O/${TAG}_bench_reference.json
timeout 300 python bench.py --no-graph --no-cpu --no-stages --steps 3 --warmup 3 > $O/${TAG}_plain.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 200 --csv --log-file $O/${TAG}_launches_bench.csv \
   python bench.py --no-graph --no-cpu --no-stages --steps 3 --warmup 3 > $O/${TAG}_ncu_launches.log 2>&1
tail -3 $O/${TAG}_tests.log 2>/dev/null; head -c 1500 $O/${TAG}_bench.json; echo; head -c 600 $O/${TAG}_bench_reference.json
