#!/bin/bash
# fused trunk vs layer-by-layer bench + ncu launch list of the fused step
mkdir -p gpurun_out
WHEATLEY_B200_TRUNK=layers bash tools/gpu_bench.sh r2f_layers --no-cpu-baseline
bash tools/gpu_launchlist.sh r2f
python tools/ncu_summary.py launches gpurun_out/r2f_launches.csv > gpurun_out/r2f_launches_summary.txt 2>&1
head -45 gpurun_out/r2f_launches_summary.txt
