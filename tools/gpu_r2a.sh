#!/bin/bash
# Round-2 first visit: full GPU parity suite (new golden / bench-shape tests), smoke, bench line.
TAG=${1:-r2a}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x -s > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${TAG}_pytest.log
grep -a "bench-shape trunk" gpurun_out/${TAG}_pytest.log
tail -5 gpurun_out/${TAG}_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/${TAG}_bench.json')); print(round(d['ms_per_step'],3), 'e2e', d['e2e']['ms_per_step'], d['e2e']['ms_per_step_mean'], 'launches', d['gpu_launches'], 'frac', d['roofline']['frac']); print(d['roofline']['share_of_step']); print(d.get('cpu_baseline'))"
tail -3 gpurun_out/${TAG}_bench.err
