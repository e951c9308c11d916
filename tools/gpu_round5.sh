#!/bin/bash
# Round step: the new GPU tests first (progressive passes, refit), then the A/B of the parking variants,
# then the whole GPU suite.
set -u
mkdir -p gpurun_out
export B200PT_QUIET=1
timeout 600 python -m pytest tests/test_progressive.py tests/test_refit.py -m gpu -q > gpurun_out/pytest_new.log 2>&1; echo "pytest new rc=$?" | tee -a gpurun_out/pytest_new.log
tail -25 gpurun_out/pytest_new.log
bash tools/ab2.sh
timeout 900 python tools/fuse_ab.py > gpurun_out/fuse_ab.log 2>&1; echo "fuse_ab rc=$?"; cat gpurun_out/fuse_ab.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
