#!/bin/bash
# usage: gpu_win3.sh TAG  -> GPU suite + bench lines of the workloads the sparse kernels matter for
TAG=${1:-w3}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/${TAG}_pytest.log
run() {   # name, env assignments..., -- bench args
  name=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 600 python bench.py --no-cpu-baseline "$@" > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err; echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/${TAG}_${name}.json'))
    r=d['roofline']
    print('  ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'], {k:round(v['avg_ms'],4) for k,v in r['kernels'].items()}, {k:round(v,3) for k,v in r['share_of_step'].items() if k.startswith('conv')})
except Exception as e:
    print('  no line', e)
PY
}
run t100 X=1 -- --workload taillard100x20
run psp272_occ2 X=1 -- --workload psp272
run psp272_occ3 B200GAT_WIN_OCC=3 -- --workload psp272
run psp272_nowin B200GAT_WIN=0 -- --workload psp272
run pspnode X=1 -- --workload psp_taillard100x20 --conflicts node
run pspnode_occ3 B200GAT_WIN_OCC=3 -- --workload psp_taillard100x20 --conflicts node
run pspclique X=1 -- --workload psp_taillard100x20 --conflicts clique
