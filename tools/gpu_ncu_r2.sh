#!/bin/bash
# ncu --set full of one layer's conv kernels (forward of the last layer, backward of the last layer) in a warm eager
# step: per step 6 x (k_fwd_win, k_clq_fwd) then 6 x (k_bwd_dst_win, k_bwd_src_win, k_clq_bwd_fused) = 30 conv
# launches (the split hubs are merged inside the *_win kernels); skip 2 steps + 10 launches, capture 5.
TAG=${1:-r2s}
mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/${TAG}_plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:'k_clq_fwd|k_fwd_fast|k_fwd_win|k_hub_merge_fwd|k_clq_bwd_fused|k_clq_bwd_mma|k_clq_bwd_elem|k_bwd_dst_fast|k_bwd_dst_win|k_bwd_src_fast|k_bwd_src_win|k_hub_merge_bwd' \
  --launch-skip 70 --launch-count 5 -o gpurun_out/${TAG}_conv_prof -f \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/${TAG}_ncu_full2.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/${TAG}_conv_prof.ncu-rep
