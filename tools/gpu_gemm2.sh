#!/bin/bash
TAG=${1:-g4}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py -q -m gpu -x > gpurun_out/${TAG}_gemm_pytest.log 2>&1; echo "gemm pytest rc=$?"
grep -v "^E  " gpurun_out/${TAG}_gemm_pytest.log | tail -15
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/${TAG}_bench.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['gpu_launches'], d['roofline']['share_of_step'])"
tail -3 gpurun_out/${TAG}_bench.err
B200GAT_GEMM_BN=128 timeout 600 python bench.py --no-cpu-baseline > gpurun_out/${TAG}_bench_bn64.json 2>> gpurun_out/${TAG}_bench.err; echo "bench bn64 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/${TAG}_bench_bn64.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['gpu_launches'], d['roofline']['share_of_step'])"
timeout 600 python tools/prof_e2e.py > gpurun_out/${TAG}_prof_e2e.log 2>&1; echo "prof rc=$?"; timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -3
grep -E "ms per eager|host:" gpurun_out/${TAG}_prof_e2e.log
