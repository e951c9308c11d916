#!/bin/bash
set -u
mkdir -p gpurun_out
export B200PT_QUIET=1
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -v "^ - \|^Parsing\|^Image" gpurun_out/pytest_gpu.log | tail -5
timeout 900 python tools/trirec_ab.py > gpurun_out/trirec_ab2.log 2>&1; echo "trirec rc=$?"
