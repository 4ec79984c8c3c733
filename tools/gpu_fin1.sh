#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/f1_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/f1_pytest.log
for bn in 0 128 0 128; do
  B200GAT_GEMM_BN=$bn timeout 600 python bench.py --no-cpu-baseline > gpurun_out/f1_bn${bn}.json 2> gpurun_out/f1_bn${bn}.err; echo "bn=$bn rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/f1_bn${bn}.json'))
print('  ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'fp32', round(d['fp32_dense']['ms_per_step'],3))
PY
done
B200GAT_GEMM_BN=128 timeout 600 python bench.py --workload taillard30x20 --no-cpu-baseline > gpurun_out/f1_30_bn128.json 2>/dev/null
timeout 600 python bench.py --workload taillard30x20 --no-cpu-baseline > gpurun_out/f1_30_bn0.json 2>/dev/null
python - <<PY
import json
for t in ('bn0','bn128'):
    d=json.load(open('gpurun_out/f1_30_%s.json'%t)); print('30x20',t, round(d['ms_per_step'],3), round(d['e2e']['ms_per_step'],3))
PY
