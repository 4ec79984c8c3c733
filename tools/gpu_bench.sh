#!/bin/bash
# One bench line (+ optional args) into gpurun_out/<tag>_bench.json
TAG=${1:-b}; shift
mkdir -p gpurun_out
timeout 900 python bench.py "$@" > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
tail -5 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.load(open('gpurun_out/${TAG}_bench.json'))
e=d['e2e']
print('ms/step', round(d['ms_per_step'],3), 'value', '%.3e'%d['value'], '| e2e mean', round(e['ms_per_step'],3), 'median', round(e.get('ms_per_step_median',0),3), e['ms_per_step_spread'], 'serial', round(e.get('ms_per_step_no_prefetch',0),3), '| launches', d['gpu_launches'])
r=d.get('roofline') or {}
print('roofline', r.get('kernel'), r.get('frac'), {k:round(v['avg_ms'],4) for k,v in (r.get('kernels') or {}).items()})
print('shares', {k:round(v,3) for k,v in (r.get('share_of_step') or {}).items()})
print('ppo', d.get('ppo',{}).get('env_steps_per_sec'), 'rollout ms', d.get('ppo',{}).get('rollout_ms_per_step'))
print('cpu', d.get('cpu_baseline'))
PY
