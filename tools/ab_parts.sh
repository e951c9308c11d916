#!/bin/bash
# A/B of variant libraries in the product configuration (no per-kernel events, B200PT_PARTS = 1..4)
mkdir -p gpurun_out
: > gpurun_out/ab_parts.log
for lib in owl_path_tracer_b200/lib/libb200pt.so owl_path_tracer_b200/lib/variants/libb200pt_*.so; do
  echo "== $(basename $lib .so)" >> gpurun_out/ab_parts.log
  B200PT_LIB=$PWD/$lib timeout 300 python tools/gpu_check.py parts 2>&1 | grep '"spp": 32' >> gpurun_out/ab_parts.log
done
python - <<'PY'
import json
for l in open('gpurun_out/ab_parts.log'):
    l=l.strip()
    if l.startswith('=='): print(l); continue
    try:
        d=json.loads(l); print('   %-7s parts %d  %.1f ms  %.0f Mrays/s'%(d['scene'],d['parts'],d['render_ms'],d['mrays']))
    except Exception: pass
PY
