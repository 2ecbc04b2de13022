#!/bin/bash
# one bench line on N GPUs -> gpurun_out/TAG_bench_Ngpu[_SORT].json   (usage: r4_multi.sh N TAG [exchange|local|replicated] [extra bench flags])
N=$1; TAG=${2:-r4x}; SORT=${3:-exchange}; shift 3
O=gpurun_out; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
F=$O/${TAG}_bench_${N}gpu_${SORT}
timeout 600 $RUN bench.py --gpus $N --steps 20 --warmup 5 --time-n1 $( [ "$SORT" = auto ] || echo --shard-sort $SORT ) "$@" 2> $F.err | grep -v "^NCCL" | tail -1 > $F.json; echo "bench rc=${PIPESTATUS[0]}"
grep "bench +" $F.err | tail -3
python - <<PY
import json
d=json.load(open("$F.json"))
print(d["n_gpus"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["parity_check"]["rows_equal_single_gpu"], d["n1_same_workload"])
print({k:round(v["ms"],2) for k,v in d["stages"].items()})
w=d.get("weak_b")
if w: print("weak_b", w["ms_per_step"], w["parity_check"]["rows_equal_single_gpu"], w["stages_ms"])
PY
