#!/bin/bash
# bench line on N GPUs (default mode) -> gpurun_out/TAG_bench_Ngpu.json ; optional NCCL parity tests first (N = 2)
N=$1; TAG=${2:-r3x}; TESTS=$3
O=gpurun_out; mkdir -p $O
if [ "$TESTS" = "tests" ]; then
  timeout 1200 python -m pytest tests/test_sharded_nccl_gpu.py -q -m gpu 2>&1 | tail -6 | tee $O/${TAG}_nccl_tests_${N}gpu.log
fi
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $RUN bench.py --gpus $N --steps 20 --warmup 5 2> $O/${TAG}_bench_${N}gpu.err | grep -v "^NCCL" | tail -1 > $O/${TAG}_bench_${N}gpu.json; echo "bench rc=${PIPESTATUS[0]}"
timeout 900 $RUN bench.py --impl reference --gpus $N --steps 3 --warmup 1 2>> $O/${TAG}_bench_${N}gpu.err | grep -v "^NCCL" | tail -1 > $O/${TAG}_bench_reference_${N}gpu.json
python - <<PY
import json
d=json.load(open("$O/${TAG}_bench_${N}gpu.json"))
print(d["n_gpus"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["parity_check"]["rows_equal_single_gpu"], d["n1_same_workload"])
print({k:round(v["ms"],2) for k,v in d["stages"].items()})
w=d["weak_b"]; print(w["ms_per_step"], w["parity_check"]["rows_equal_single_gpu"])
r=json.load(open("$O/${TAG}_bench_reference_${N}gpu.json")); print("reference", r["value"], r["cpu_baseline"]["cores"], r["cpu_baseline"]["sample"][:80])
PY
