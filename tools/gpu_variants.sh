#!/bin/bash
TAG=${1:-g9}
mkdir -p gpurun_out
show() { python -c "
import json,sys; d=json.load(open('$1')); s=d['roofline']['share_of_step']; print('$2', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],2), {k: round(v*6.3,3) for k,v in s.items()})"; }
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "rc=$?"; show gpurun_out/${TAG}_bench.json default
for m in 3 4; do
  B200GAT_LIB=$PWD/tools/variants/libb200gat_minb$m.so timeout 300 python bench.py --no-cpu-baseline > gpurun_out/${TAG}_bench_minb$m.json 2>> gpurun_out/${TAG}_bench.err; echo "rc=$?"; show gpurun_out/${TAG}_bench_minb$m.json minb$m
done
tail -3 gpurun_out/${TAG}_bench.err
