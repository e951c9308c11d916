#!/bin/bash
# round-2 GPU pass R: pipeline slots x start fraction at the headline 1024 spp, after the builder change
export B200PT_QUIET=1
for cfg in "3 0.5" "3 0.35" "3 0.65" "4 0.5" "4 0.7" "2 0.5"; do
  set -- $cfg
  timeout 600 python bench.py --steps 6 --warmup 2 --pipeline $1 --start-fraction $2 --streams-mode 0 --no-cpu-baseline --no-parity-check > gpurun_out/r02r_p$1_f$2.json 2> gpurun_out/r02r.err || { echo "failed $cfg"; tail -3 gpurun_out/r02r.err; }
  python - <<PY
import json
d=json.load(open('gpurun_out/r02r_p$1_f$2.json')); print('slots $1 start $2:', round(d['value'],1), 'Mrays/s', round(d['ms_per_step'],1), 'ms; e2e', round(d['e2e']['value'],1), 'equal', d['frames']['pipelined_frame_equals_blocking_frame'])
PY
done
