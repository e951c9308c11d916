#!/bin/bash
export B200PT_QUIET=1
V=owl_path_tracer_b200/lib/variants
timeout 900 python tools/ab.py --spp 32 --timing --check base=owl_path_tracer_b200/lib/libb200pt.so ploc4=$V/libb200pt_ploc4.so ploc16=$V/libb200pt_ploc16.so ploc32=$V/libb200pt_ploc32.so cost06=$V/libb200pt_cost06.so base2=owl_path_tracer_b200/lib/libb200pt.so > gpurun_out/r02i_ab.jsonl 2>&1; cut -c1-420 gpurun_out/r02i_ab.jsonl | grep -v check
timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/r02i_bench.json 2> gpurun_out/r02i_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02i_bench.json')); print({k:d[k] for k in ('value','ms_per_step','frames','parity_check','e2e','throughput_mode','kernel_ms_per_step')}); print(d['roofline'])
PY
