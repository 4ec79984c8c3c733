#!/bin/bash
# 2-GPU runs: weak scaling of the headline workload, strong scaling of BASELINE config 5 (30x20, 128 graphs)
mkdir -p gpurun_out
N=${1:-2}
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --no-cpu-baseline "$@" > gpurun_out/${tag}.json 2> gpurun_out/${tag}.err; echo "$tag rc=$?"; tail -2 gpurun_out/${tag}.err; python - <<PY
import json
try:
    d=json.load(open('gpurun_out/${tag}.json')); e=d['e2e']
    print('${tag}', 'ms/step', round(d['ms_per_step'],3), 'value %.3e'%d['value'], 'e2e mean', round(e['ms_per_step'],3), 'scaling', d['scaling'], d['config'].get('graphs_per_gpu'))
except Exception as ex: print('no json', ex)
PY
}
run r2z_weak_n$N
run r2z_strong30x20_n$N --scaling strong --workload taillard30x20 --graphs 128
