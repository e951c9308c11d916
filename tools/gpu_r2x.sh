#!/bin/bash
# round-2 GPU pass X: last check of the tree as committed -- GPU suite, smoke, bench line with the driver's arguments
export B200PT_QUIET=1
timeout 1800 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02x_bench.json 2> gpurun_out/r02x_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02x_bench.json')); print({k:d[k] for k in ('value','ms_per_step','e2e','kernel_ms_per_step','gpu_launches','clocks')}); print(d['frames']['single_frame_mrays'], d['parity_check']['bit_equal'], d['roofline']['frac'])
PY
