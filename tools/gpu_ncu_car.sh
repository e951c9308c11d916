#!/bin/bash
# ncu capture of k_shade on the car scene (glass / clearcoat / anisotropic metal: the Disney lobes' binary64 math)
set -u
mkdir -p gpurun_out
export B200PT_QUIET=1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 400 -c 2 -f -o gpurun_out/shade_car \
    python tools/gpu_check.py ab > gpurun_out/ncu_shade_car.log 2>&1
echo "car shade capture rc=$?"; ls -la gpurun_out/shade_car.ncu-rep
