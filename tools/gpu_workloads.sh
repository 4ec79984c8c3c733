#!/bin/bash
# bench lines for the other BASELINE.json configurations (graph shapes C1, C2, C5 of SURVEY 8d)
mkdir -p gpurun_out
for w in jssp10x10 taillard15x15 taillard30x20; do
  timeout 300 python bench.py --workload $w --steps 10 --warmup 3 --cpu-seconds 6 > gpurun_out/wl_$w.json 2> gpurun_out/wl_$w.err; echo "$w rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/wl_$w.json')); print('$w', 'graphs', d['config']['graphs_per_gpu'], 'edges', d['config']['edges_per_gpu'], 'ms', round(d['ms_per_step'],3), 'value', '%.3g'%d['value'], 'e2e', '%.3g'%d['e2e']['value'], 'cpu', '%.3g'%d['cpu_baseline']['value'], 'ppo', round(d['ppo']['env_steps_per_sec'],1))"
  tail -2 gpurun_out/wl_$w.err
done
