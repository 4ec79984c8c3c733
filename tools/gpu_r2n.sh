#!/bin/bash
bash tools/gpu_bench.sh r2n_psp272 --workload psp272
bash tools/gpu_bench.sh r2n_psp100x20 --workload psp_taillard100x20
bash tools/gpu_bench.sh r2n_psp100x20_clique --workload psp_taillard100x20 --conflicts clique --no-cpu-baseline
bash tools/gpu_bench.sh r2n --no-cpu-baseline
