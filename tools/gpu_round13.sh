#!/bin/bash
set -u
mkdir -p gpurun_out
export B200PT_QUIET=1
timeout 900 python tools/trirec_ab.py > gpurun_out/trirec_ab.log 2>&1; echo "trirec rc=$?"; cat gpurun_out/trirec_ab.log | cut -c1-330
