#!/bin/bash
# round-2 GPU pass T (pooled triangle phase): final evidence on one B200 — GPU suite, bench (driver's arguments), reference arm, out-of-L2 line, ncu
export B200PT_QUIET=1
timeout 1800 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02t_bench.json 2> gpurun_out/r02t_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --gpus 1 --steps 2 --warmup 1 > gpurun_out/r02t_ref.json 2> gpurun_out/r02t_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --detail 7200000 --steps 4 --warmup 2 --streams-mode 0 --no-cpu-baseline > gpurun_out/r02t_bench_detail7p2M.json 2> gpurun_out/r02t_detail.err; echo "detail rc=$?"
python - <<'PY'
import json
for f in ('r02t_bench','r02t_bench_detail7p2M'):
    d=json.load(open(f'gpurun_out/{f}.json')); print(f, {k:d[k] for k in ('value','ms_per_step','frames','parity_check','e2e','kernel_ms_per_step')}); print(d['roofline'])
PY
bash tools/gpu_ncu.sh 2>&1 | tail -4
