#!/bin/bash
export B200PT_QUIET=1
V=owl_path_tracer_b200/lib/variants
timeout 600 python tools/ab.py --spp 64 --scenes dragon base=owl_path_tracer_b200/lib/libb200pt.so park2=$V/libb200pt_park2.so park5=$V/libb200pt_park5.so park8=$V/libb200pt_park8.so base2=owl_path_tracer_b200/lib/libb200pt.so > gpurun_out/r02j_park_ab.jsonl 2>&1; cut -c1-200 gpurun_out/r02j_park_ab.jsonl
B="python bench.py --steps 6 --warmup 2 --streams-mode 0 --no-cpu-baseline --no-parity-check --spp 256 --lib $V/libb200pt_dev.so"
for cfg in "2 2 0.5" "1 2 0.5" "1 3 0.5" "1 3 0.7" "2 3 0.6" "1 4 0.7"; do set -- $cfg
  B200PT_PARTS=$1 timeout 300 $B --pipeline $2 --start-fraction $3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('parts $1 slots $2 start $3 ->', round(d['value'],1), 'Mrays/s; single frame', round(d['frames']['single_frame_mrays'],1))"
done | tee gpurun_out/r02j_pipeline_parts.txt
