#!/bin/bash
# quick GPU visit: parity tests + dragon/car timing of the main library
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/gpu_check.py ab > gpurun_out/ab_main.log 2>&1; cat gpurun_out/ab_main.log
[ -d owl_path_tracer_b200/lib/variants ] && bash tools/ab.sh
