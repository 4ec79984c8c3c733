#!/bin/bash
# ncu --set full of 14 consecutive tcgen05 GEMM / split-K launches of a warm eager step (forward MLP + projection of a
# middle layer, then part of the backward), TF32 mode, for profiles/<tag>_gemm_kernels_full.txt
TAG=${1:-r2z}
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_gemm_tf32|k_splitk_reduce' \
  --launch-skip ${2:-230} --launch-count 14 -o gpurun_out/${TAG}_gemm_prof -f \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/${TAG}_ncu_gemm.log 2>&1; echo "ncu gemm rc=$?"
python tools/ncu_summary.py full gpurun_out/${TAG}_gemm_prof.ncu-rep > gpurun_out/${TAG}_gemm_kernels_full.txt 2>&1
rm -f gpurun_out/${TAG}_gemm_prof.ncu-rep
cut -c1-230 gpurun_out/${TAG}_gemm_kernels_full.txt | head -12
