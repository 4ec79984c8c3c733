#!/bin/bash
# ncu --set full of one layer's conv kernels, forward and backward (for roofline.traffic): the launches of the
# last layer of the third warm step (2 steps x 48 launches + 5 forward layers x 3 skipped).
TAG=${1:-r1m}
mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/${TAG}_plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:'k_clq_fwd|k_fwd_fast|k_hub_merge_fwd|k_clq_bwd_mma|k_clq_bwd_elem|k_bwd_dst_fast|k_bwd_src_fast|k_hub_merge_bwd' \
  --launch-skip 111 --launch-count 8 -o gpurun_out/${TAG}_conv_prof -f \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/${TAG}_ncu_full2.log 2>&1; echo "ncu full rc=$?"
