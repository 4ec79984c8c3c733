#!/bin/bash
TAG=${1:-g6}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
timeout 600 python bench.py --gemm cublas --no-cpu-baseline > gpurun_out/${TAG}_bench_cublas.json 2>> gpurun_out/${TAG}_bench.err; echo "bench cublas rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/${TAG}_bench_cublas.json')); print('cublas', d['ms_per_step'], d['e2e']['ms_per_step'], d['gpu_launches'])"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_gemm_tf32|k_splitk' \
  --launch-skip 80 --launch-count 14 -o gpurun_out/${TAG}_gemm_prof -f \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
