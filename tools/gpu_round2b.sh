#!/bin/bash
# Round-2 final measurement visit (one GPU): ncu --set full of the conv kernels -> DRAM traffic file keyed to the
# kernel sources, then the bench lines of every workload (both arms), the launch list of the default step, smoke.
TAG=${1:-r2z}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/${TAG}_smoke.log
bash tools/gpu_ncu_r2.sh ${TAG}
python tools/ncu_summary.py full gpurun_out/${TAG}_conv_prof.ncu-rep > gpurun_out/${TAG}_conv_kernels_full.txt 2>&1
cp gpurun_out/${TAG}_conv_kernels_full.txt profiles/r2z_conv_kernels_full.txt
python tools/make_traffic.py profiles/r2z_conv_kernels_full.txt profiles/r2_traffic.json taillard100x20 8 64 4 > /dev/null
cp profiles/r2_traffic.json gpurun_out/${TAG}_traffic.json
rm -f gpurun_out/${TAG}_conv_prof.ncu-rep      # 50 MB: the summary is what is kept
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
        print(f.split('${TAG}_')[1], 'ms', round(d['ms_per_step'],3), 'value %.3e'%d['value'], 'e2e', round(e.get('ms_per_step',0) or 0,3), 'fp32', (d.get('fp32_dense') or {}).get('ms_per_step'), 'frac', r.get('frac'), 'traffic', r.get('traffic'), 'ppo', (d.get('ppo') or {}).get('env_steps_per_sec'), 'L', d.get('gpu_launches'))
    except Exception as ex: print(f, 'ERR', ex)
PY
bash tools/gpu_launchlist.sh ${TAG}
python tools/ncu_summary.py launches gpurun_out/${TAG}_launches.csv > gpurun_out/${TAG}_launches_summary.txt 2>&1; head -24 gpurun_out/${TAG}_launches_summary.txt
gzip -f gpurun_out/${TAG}_launches.csv
