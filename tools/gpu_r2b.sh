#!/bin/bash
# round-2 GPU pass B2: A/B of walker / shade variants, timeline of the base walker
export B200PT_QUIET=1
V=owl_path_tracer_b200/lib/variants
args="base=owl_path_tracer_b200/lib/libb200pt.so"
for n in smem8 smem12 smem16 libm smem12libm shade5 leaf2v2 leaf2v3 leaf2v4 leaf4v2 leaf4v3 leaf4v4 leaf6v2 leaf6v4; do args="$args $n=$V/libb200pt_$n.so"; done
timeout 1500 python tools/ab.py --spp 32 --timing --check $args base2=owl_path_tracer_b200/lib/libb200pt.so > gpurun_out/r02b_ab.jsonl 2>&1
cat gpurun_out/r02b_ab.jsonl | cut -c1-400
timeout 600 python tools/timeline.py $V/libb200pt_timeline.so > gpurun_out/r02b_timeline_base.jsonl 2>&1; cut -c1-600 gpurun_out/r02b_timeline_base.jsonl | tail -8
