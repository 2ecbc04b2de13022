#!/bin/bash
# Multi-GPU evidence on N GPUs of one box: the bench line (configs[4] strong scaling + weak configs[1] +
# parity_check), the reference arm as the driver launches it, and the balanced split on configs[3].
#   gpurun --gpus N -- bash tools/r2_multi.sh N TAG
N=$1; TAG=${2:-r2x}
O=gpurun_out; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $RUN bench.py --gpus $N --steps 20 --warmup 5 2> $O/${TAG}_bench_${N}gpu.err | grep -v "^NCCL" | tail -1 > $O/${TAG}_bench_${N}gpu.json; echo "bench rc=${PIPESTATUS[0]}"
timeout 600 $RUN tools/balance_bench.py --scene configs3 --spheres 16777216 --frames 5 --mode replicate_build 2>&1 | grep -E "^balance" > $O/${TAG}_balance_configs3_${N}gpu.txt
head -c 600 $O/${TAG}_bench_${N}gpu.json; echo; cat $O/${TAG}_balance_configs3_${N}gpu.txt
