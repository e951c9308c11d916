#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/gpu_check.py > gpurun_out/check.log 2>&1; echo "check rc=$?"; cat gpurun_out/check.log | cut -c1-600
bash tools/ab.sh
