#!/bin/bash
set -u
mkdir -p gpurun_out
export B200PT_QUIET=1
timeout 600 python tools/refit_debug.py > gpurun_out/refit_debug.log 2>&1; echo "refit_debug rc=$?"; tail -20 gpurun_out/refit_debug.log
timeout 600 python -m pytest tests/test_progressive.py tests/test_refit.py -m gpu -q > gpurun_out/pytest_new.log 2>&1; echo "pytest new rc=$?" | tee -a gpurun_out/pytest_new.log
grep -v "^ - \|^Parsing\|^Image" gpurun_out/pytest_new.log | tail -25
bash tools/ab2.sh
