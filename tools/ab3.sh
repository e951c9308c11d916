#!/bin/bash
# A/B of variant libraries with per-kernel times: tools/gpu_check.py ab (one stream, events) and ab2 (product configuration)
mkdir -p gpurun_out
: > gpurun_out/ab3.log
for lib in owl_path_tracer_b200/lib/variants/libb200pt_*.so; do
  echo "== $(basename $lib .so)" >> gpurun_out/ab3.log
  B200PT_LIB=$PWD/$lib timeout 300 python tools/gpu_check.py ab 2>&1 | grep "^{" >> gpurun_out/ab3.log
  B200PT_LIB=$PWD/$lib timeout 300 python tools/gpu_check.py ab2 2>&1 | grep "^{" >> gpurun_out/ab3.log
done
python - <<'PY'
import json
for l in open('gpurun_out/ab3.log'):
    if l.startswith('=='): print(l.strip()); continue
    d=json.loads(l)
    if 'primary_mismatch' in d: print('   parity', d['mean_bit_mismatch'], d['rays_gpu']==d['rays_oracle']); continue
    print('   %s %d spp: %.1f ms ext %.1f shade %.1f  %.0f Mrays/s'%(d['scene'], d['spp'], d['render_ms'], d['extend_ms'], d['shade_ms'], d['mrays']))
PY
