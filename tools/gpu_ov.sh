#!/bin/bash
TAG=${1:-ov}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/${TAG}_pytest.log
for ov in 1 0; do
for wl in taillard100x20 psp272 taillard30x20; do
  B200GAT_DW_OVERLAP=$ov timeout 600 python bench.py --workload $wl --no-cpu-baseline > gpurun_out/${TAG}${ov}_${wl}.json 2> gpurun_out/${TAG}${ov}_${wl}.err; echo "overlap=$ov $wl rc=$?"
  tail -2 gpurun_out/${TAG}${ov}_${wl}.err
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/${TAG}${ov}_${wl}.json'))
    print('  ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'serial', round(d['e2e']['ms_per_step_no_prefetch'],3), 'fp32', round(d['fp32_dense']['ms_per_step'],3), 'ppo', round(d['ppo']['env_steps_per_sec']))
except Exception as e:
    print('  no line', e)
PY
done
done
