#!/bin/bash
# GPU visit for the tcgen05 dense path: GEMM parity tests first (bounded), then the whole suite, the bench and
# an ncu launch list of the eager step.
TAG=${1:-g1}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py -q -m gpu -x > gpurun_out/${TAG}_gemm_pytest.log 2>&1; echo "gemm pytest rc=$?"
grep -v "^E  " gpurun_out/${TAG}_gemm_pytest.log | tail -15
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
cat gpurun_out/${TAG}_bench.json; tail -5 gpurun_out/${TAG}_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
