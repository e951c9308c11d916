#!/bin/bash
# A/B timing of variant libraries on the GPU box (dragon + car, 1080p, 32 spp, depth 8)
mkdir -p gpurun_out
: > gpurun_out/ab.log
for lib in owl_path_tracer_b200/lib/variants/libb200pt_*.so; do
  name=$(basename $lib .so)
  echo "== $name" >> gpurun_out/ab.log
  B200PT_LIB=$PWD/$lib timeout 300 python tools/gpu_check.py ab >> gpurun_out/ab.log 2>&1
done
cat gpurun_out/ab.log
