#!/bin/bash
# GPU visit for the NEE mode: all parity tests, NEE timing/convergence, main-path timing
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python tools/gpu_check.py nee > gpurun_out/nee.log 2>&1; cat gpurun_out/nee.log
timeout 300 python tools/gpu_check.py ab > gpurun_out/ab_main.log 2>&1; cat gpurun_out/ab_main.log
