#!/bin/bash
# Windowed sparse kernels: GPU suite, then A/B bench lines (B200GAT_WIN=0 / default / 3 CTAs per SM) on two workloads
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/win_pytest.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/win_pytest.log
for v in "0 2" "-1 2" "-1 3"; do
  set -- $v
  tag="win${1}_occ${2}"
  for wl in taillard100x20 psp272; do
    if [ "$1" = "-1" ]; then unset B200GAT_WIN; else export B200GAT_WIN=$1; fi
    B200GAT_WIN_OCC=$2 timeout 600 python bench.py --workload $wl --no-cpu-baseline > gpurun_out/${tag}_${wl}.json 2> gpurun_out/${tag}_${wl}.err; echo "$tag $wl rc=$?"
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/${tag}_${wl}.json'))
    r=d['roofline']
    print('  ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'], {k:round(v['avg_ms'],4) for k,v in r['kernels'].items()}, {k:round(v,3) for k,v in r['share_of_step'].items()})
except Exception as e:
    print('  no line', e)
PY
  done
done
