#!/bin/bash
# round-2 GPU pass W: persistent grid of the walker at 6 (product) / 5 / 4 blocks per SM under the frame pipeline
# (fewer walker blocks leave registers for k_shade blocks of the other frames to be co-resident)
export B200PT_QUIET=1
L=owl_path_tracer_b200/lib
for v in "" variants/libb200pt_g5.so variants/libb200pt_g4.so; do
  if [ -z "$v" ]; then LIBARG=""; name=base; else LIBARG="--lib $L/$v"; name=$(basename $v .so); fi
  timeout 600 python bench.py --steps 6 --warmup 2 --streams-mode 0 --no-cpu-baseline --no-parity-check $LIBARG > gpurun_out/r02w_$name.json 2> gpurun_out/r02w.err || { echo "failed $name"; tail -3 gpurun_out/r02w.err; }
  python - <<PY
import json
d=json.load(open('gpurun_out/r02w_$name.json')); print('$name:', round(d['value'],1), 'Mrays/s', round(d['ms_per_step'],1), 'ms; blocking frame', round(d['frames']['single_frame_ms'],1), 'ms; kernel', {k: round(v,1) for k,v in d['kernel_ms_per_step'].items()})
PY
done
