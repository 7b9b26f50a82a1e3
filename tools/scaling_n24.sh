#!/bin/bash
# 2- and 4-GPU points of the C2 scaling row (run on a 4-GPU box)
out=gpurun_out/scaling_n24.jsonl
: > $out
port=29720
for n in 2 4; do
  port=$((port+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $n --no-cpu-baseline --steps 3 --warmup 3 2>gpurun_out/scaling_err.log | grep '^{' | tail -1 >> $out
done
python - <<'PY'
import json
for l in open("gpurun_out/scaling_n24.jsonl"):
    d = json.loads(l)
    print(d["config"]["workload"][:24], "N", d["n_gpus"], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms", round(d["ms_per_step"], 2), d["config"]["tracer"], d["config"]["megakernel_form"])
PY
