#!/bin/bash
# round-2 GPU pass K: whole GPU suite, smoke, bench (pipeline of 3 one-part contexts)
export B200PT_QUIET=1
timeout 1800 python -m pytest tests -x -q -m gpu --durations=8 2>&1 | tail -16
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02k_bench.json 2> gpurun_out/r02k_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02k_bench.json')); print({k:d[k] for k in ('value','ms_per_step','frames','parity_check','e2e','throughput_mode','kernel_ms_per_step','gpu_launches')})
PY
