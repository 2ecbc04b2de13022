#!/bin/bash
# One `ncu --set full` capture of the pipeline kernels of bench.py (one GPU, after bench.py exited 0 without ncu).
TAG=${1:-rX}
O=gpurun_out
mkdir -p $O
timeout 300 python bench.py --no-graph --steps 2 --warmup 3 > /dev/null 2>&1 || exit 1
timeout 1200 ncu --set full --clock-control none --import-source on \
  -k regex:"candidates|confirm|emit|build_refit|reduce_kernel|morton|onesweep|scan_kernel|histogram" \
  --launch-skip 40 -c 14 -o $O/${TAG}_full -f python bench.py --no-graph --steps 2 --warmup 3 > $O/${TAG}_ncu_full.log 2>&1
