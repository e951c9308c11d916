#!/bin/bash
set -u
mkdir -p gpurun_out
export B200PT_QUIET=1
timeout 600 python tools/l2_ab.py > gpurun_out/l2_ab.log 2>&1; echo "l2_ab rc=$?"; cat gpurun_out/l2_ab.log
bash tools/ab2.sh | grep "^==\|1080" | cut -c1-260
