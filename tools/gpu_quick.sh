#!/bin/bash
# tests + bench + host profile of the eager step
TAG=${1:-q}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -2
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/${TAG}_bench.json')); s=d['roofline']['share_of_step']; print(round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],2), 'launches', d['gpu_launches'])"
tail -2 gpurun_out/${TAG}_bench.err
timeout 300 python tools/prof_e2e.py > gpurun_out/${TAG}_prof.log 2>&1; head -45 gpurun_out/${TAG}_prof.log | cut -c1-150
