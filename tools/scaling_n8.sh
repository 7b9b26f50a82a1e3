#!/bin/bash
# 8-GPU points of the scaling table after a kernel change (the N=1 values come from the single-GPU bench lines)
out=gpurun_out/scaling_n8.jsonl
: > $out
port=29700
for args in "--steps 3 --warmup 3" "--steps 2 --warmup 3 --workload c5"; do
  port=$((port+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus 8 --no-cpu-baseline $args 2>gpurun_out/scaling_err.log | grep '^{' | tail -1 >> $out
done
python - <<'PY'
import json
for l in open("gpurun_out/scaling_n8.jsonl"):
    d = json.loads(l)
    print(d["config"]["workload"][:24], "N", d["n_gpus"], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms", round(d["ms_per_step"], 2), d["config"]["tracer"], d["config"]["megakernel_form"])
PY
