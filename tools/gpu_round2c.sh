#!/bin/bash
# Bench lines of every workload (both arms) with the current bench.py; no profiler.
TAG=${1:-r2z}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${TAG}_pytest.log
b() { t=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/${TAG}_$t.json 2> gpurun_out/${TAG}_$t.err; echo "$t rc=$?"; tail -c 300 gpurun_out/${TAG}_$t.err | tail -2; }
b bench
b bench_ref --impl reference --steps 5 --warmup 1
b bench_psp272 --workload psp272
b bench_psp272_ref --workload psp272 --impl reference --steps 20 --warmup 2
b bench_psp100x20 --workload psp_taillard100x20 --no-cpu-baseline
b bench_psp100x20_clique --workload psp_taillard100x20 --conflicts clique --no-cpu-baseline
b bench_taillard30x20 --workload taillard30x20 --no-cpu-baseline
b bench_taillard15x15 --workload taillard15x15 --no-cpu-baseline
b bench_jssp10x10 --workload jssp10x10 --no-cpu-baseline
b bench_n1_30x20x128 --workload taillard30x20 --graphs-per-gpu 128 --no-cpu-baseline
python - <<PY
import json,glob
for f in sorted(glob.glob('gpurun_out/${TAG}_bench*.json')):
    try:
        d=json.load(open(f)); e=d.get('e2e',{}); r=d.get('roofline') or {}
        print(f.split('${TAG}_')[1], 'ms', round(d['ms_per_step'],3), 'value %.3e'%d['value'], 'e2e', round(e.get('ms_per_step',0) or 0,3), e.get('ms_per_step_by_prefetch_placement'), 'fp32', (d.get('fp32_dense') or {}).get('ms_per_step'), 'frac', r.get('frac'), 'traffic', r.get('traffic'), 'ppo', (d.get('ppo') or {}).get('env_steps_per_sec'), 'L', d.get('gpu_launches'))
    except Exception as ex: print(f, 'ERR', ex)
PY
