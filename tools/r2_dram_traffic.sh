#!/bin/bash
# DRAM traffic of the two launches bench.py quotes; run on the GPU box: bash tools/r2_dram_traffic.sh
set -e
mkdir -p gpurun_out
python tools/r2_traffic_cmd.py > gpurun_out/r2_traffic_plain.log 2>&1 &&
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,gpu__time_duration.sum --clock-control none \
    -k regex:'render_sorted_kernel|trace_persistent_kernel' --csv --log-file gpurun_out/r2_dram_traffic.csv \
    python tools/r2_traffic_cmd.py > gpurun_out/r2_traffic_ncu.log 2>&1
python tools/r2_traffic_json.py gpurun_out/r2_dram_traffic.csv gpurun_out/r2_traffic_plain.log > gpurun_out/r2_dram_traffic.json
cat gpurun_out/r2_dram_traffic.json
