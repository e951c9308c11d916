#!/bin/bash
set -u
mkdir -p gpurun_out
export B200PT_QUIET=1
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -v "^ - \|^Parsing\|^Image" gpurun_out/pytest_gpu.log | tail -40
