#!/bin/bash
TAG=${1:-r1k}
mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/${TAG}_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_clq_fwd|k_clq_bwd_mma|k_clq_bwd_elem|k_fwd_fast' \
  --launch-skip 8 --launch-count 6 -o gpurun_out/${TAG}_clq_prof -f \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
