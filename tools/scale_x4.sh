#!/bin/bash
# bench.py --density x4 at 2, 4 and 8 GPUs of one box (run under gpurun --gpus 8)
mkdir -p gpurun_out
for N in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29540 + N)) \
    bench.py --gpus $N --density x4 --steps 50 2> gpurun_out/r01_h_bench_x4_${N}gpu.err | grep '^{' > gpurun_out/r01_h_bench_x4_${N}gpu.json
  python - gpurun_out/r01_h_bench_x4_${N}gpu.json <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(d["n_gpus"], round(d["value"], 1), d["unit"], [(k["name"], round(k["ms"] / max(k["launches"], 1), 4)) for k in d.get("kernels", [])])
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
