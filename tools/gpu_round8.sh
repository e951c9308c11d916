#!/bin/bash
set -u
mkdir -p gpurun_out
export B200PT_QUIET=1
B200PT_LIB=$PWD/owl_path_tracer_b200/lib/variants/libb200pt_timeline.so timeout 600 python tools/timeline.py > gpurun_out/timeline.log 2>&1; echo "timeline rc=$?"; cat gpurun_out/timeline.log
