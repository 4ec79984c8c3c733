#!/bin/bash
# Whole GPU suite (+ optional -k filter), output in gpurun_out/<tag>_pytest.log
TAG=${1:-t}; shift
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q "$@" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${TAG}_pytest.log
tail -40 gpurun_out/${TAG}_pytest.log
