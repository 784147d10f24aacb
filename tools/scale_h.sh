#!/bin/bash
# r01_h: bit identity at 8 GPUs, then bench.py (uniform at 4 and 8, x4 at 8); run under gpurun --gpus 8
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node 8 --master-port 29520 tools/dist_check.py 163842 5 2>&1 | grep "rank" > gpurun_out/r01_h_dist_check_8gpu.log
grep -c "p2p==single True nccl==single True maxdisp equal True" gpurun_out/r01_h_dist_check_8gpu.log
run() {  # N density tag
  timeout 300 $TR --nproc-per-node $1 --master-port $((29530 + $1)) bench.py --gpus $1 --density $2 2>/dev/null | grep '^{' > gpurun_out/r01_h_bench_$3_$1gpu.json
  python - gpurun_out/r01_h_bench_$3_$1gpu.json <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(d["n_gpus"], d["config"]["density"], round(d["value"], 1), d["unit"], [(k["name"], round(k["ms"] / max(k["launches"], 1), 4)) for k in d.get("kernels", [])])
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
}
run 4 uniform uniform
run 8 uniform uniform
run 8 x4 x4b
