#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/win_pytest.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/win_pytest.log
for wl in taillard100x20 psp272; do
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/win_${wl}_launches.csv \
  python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/win_ncu_${wl}.log 2>&1; echo "ncu $wl rc=$?"
python tools/ncu_summary.py launches gpurun_out/win_${wl}_launches.csv | head -24
done
