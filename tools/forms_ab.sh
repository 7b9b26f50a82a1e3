# megakernel forms at 32 spp: SPK_RENDER_PHASED=<form> (1000 + warps*10 + CTAs/SM; 831 = phased kernel)
j='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d["ms_per_step"], d["value"])'
for f in ${C2_FORMS:-1162 1084 831}; do
  echo -n "c2 form $f: "; SPK_RENDER_PHASED=$f python bench.py --spp 32 --steps 5 --warmup 3 --tracer megakernel --no-cpu-baseline 2>/dev/null | python -c "$j"
done
for f in ${C5_FORMS:-1084 1162}; do
  echo -n "c5 form $f: "; SPK_RENDER_PHASED=$f python bench.py --workload c5 --spp 32 --steps 3 --warmup 3 --tracer megakernel --no-cpu-baseline 2>/dev/null | python -c "$j"
done
