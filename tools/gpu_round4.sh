#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
bash tools/ab.sh
timeout 900 python tools/scene_sweep.py > gpurun_out/scene_sweep.log 2>&1; echo "sweep rc=$?"; cut -c1-900 gpurun_out/scene_sweep.log
