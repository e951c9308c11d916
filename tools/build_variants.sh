#!/bin/bash
# A/B libraries of the CUDA path: tools/build_variants.sh name "-DPT_X=.. -DPT_Y=.." [name flags]...
# -> owl_path_tracer_b200/lib/variants/libb200pt_<name>.so   (bound explicitly: tools/ab.py name=<path>, owl_path_tracer_b200.use_library)
set -e
cd "$(dirname "$0")/.."
CSRC=owl_path_tracer_b200/csrc
OUT=owl_path_tracer_b200/lib/variants
mkdir -p $OUT
NVFLAGS="-std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC -Xcompiler -ffp-contract=off --expt-relaxed-constexpr -cudart static"
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  ( d=$(mktemp -d)
    for f in api wavefront bvh_build mesh_dedup; do nvcc $NVFLAGS $flags -c -o $d/$f.o $CSRC/$f.cu & done; wait
    nvcc -shared -cudart static -gencode arch=compute_100a,code=sm_100a -o $OUT/libb200pt_$name.so $d/api.o $d/wavefront.o $d/bvh_build.o $d/mesh_dedup.o
    rm -rf $d; echo "built $OUT/libb200pt_$name.so  [$flags]" ) &
done
wait
