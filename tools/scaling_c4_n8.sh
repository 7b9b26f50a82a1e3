#!/bin/bash
# C4 (city-10M at 4K, 128 of 1024 spp) on 8 GPUs
port=29741
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port \
  bench.py --gpus 8 --no-cpu-baseline --steps 2 --warmup 3 --workload c4 --spp 128 2>gpurun_out/scaling_err.log | grep '^{' | tail -1 > gpurun_out/scaling_c4_n8.jsonl
python - <<'PY'
import json
for l in open("gpurun_out/scaling_c4_n8.jsonl"):
    d = json.loads(l)
    print(d["config"]["workload"][:24], "N", d["n_gpus"], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms", round(d["ms_per_step"], 2), d["config"]["tracer"], d["config"]["megakernel_form"])
PY
