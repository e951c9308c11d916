#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
B200PT_PARTS=3 timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_p3.log 2>&1; echo "pytest(parts=3) rc=$?" | tee -a gpurun_out/pytest_gpu_p3.log
tail -5 gpurun_out/pytest_gpu_p3.log
timeout 600 python tools/gpu_check.py parts > gpurun_out/parts.log 2>&1; cat gpurun_out/parts.log
