#!/bin/bash
# 1/2/4/8-GPU runs of bench.py on one box: C2 (headline), C4 bounded to 128 spp, C5 at 1 and 8 GPUs.
# Writes one JSON line per run into gpurun_out/scaling_r1.jsonl
out=gpurun_out/scaling_r1.jsonl
: > $out
port=29600
run() { # n, args...
  n=$1; shift
  port=$((port+1))
  if [ "$n" = 1 ]; then
    python bench.py --gpus 1 "$@" 2>gpurun_out/scaling_err.log | tail -1 >> $out
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $n "$@" 2>gpurun_out/scaling_err.log | grep '^{' | tail -1 >> $out
  fi
}
for n in 1 2 4 8; do run $n --steps 3 --warmup 3 --no-cpu-baseline; done
for n in 1 2 4 8; do run $n --steps 2 --warmup 3 --no-cpu-baseline --workload c4 --spp 128; done
for n in 1 8; do run $n --steps 2 --warmup 3 --no-cpu-baseline --workload c5; done
python - <<'PY'
import json
for l in open("gpurun_out/scaling_r1.jsonl"):
    d = json.loads(l)
    print(d["config"]["workload"][:24], "N", d["n_gpus"], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms", round(d["ms_per_step"], 2), d["config"]["tracer"])
PY
