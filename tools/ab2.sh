#!/bin/bash
# A/B of variant libraries in the product configuration (tools/gpu_check.py ab2): parity at 640x360 + dragon/car 1080p 128 spp
mkdir -p gpurun_out
: > gpurun_out/ab2.log
for lib in owl_path_tracer_b200/lib/variants/libb200pt_*.so; do
  name=$(basename $lib .so)
  echo "== $name" >> gpurun_out/ab2.log
  B200PT_LIB=$PWD/$lib timeout 300 python tools/gpu_check.py ab2 2>&1 | grep -v "^$" >> gpurun_out/ab2.log
done
cat gpurun_out/ab2.log
