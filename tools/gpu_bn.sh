#!/bin/bash
mkdir -p gpurun_out
for bn in 0 128 64; do
  B200GAT_GEMM_BN=$bn timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bn${bn}.json 2> gpurun_out/bn${bn}.err; echo "bn=$bn rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/bn${bn}.json'))
r=d['roofline']
print('  ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'fp32', round(d['fp32_dense']['ms_per_step'],3), {k:round(v,3) for k,v in r['share_of_step'].items() if k.startswith('gemm')})
PY
done
