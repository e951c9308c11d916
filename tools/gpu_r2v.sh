#!/bin/bash
# round-2 GPU pass V (N GPUs, final walker): bench at N=8 launched like the driver does
export B200PT_QUIET=1
N=${1:-8}
timeout 700 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02v_bench_${N}gpu.json 2> gpurun_out/r02v_bench_${N}gpu.err; echo "bench rc=$?"; grep -v "^$\|\*\*\*\|OMP_NUM" gpurun_out/r02v_bench_${N}gpu.err | tail -4 | cut -c1-300
python - <<PY
import json
d=json.load(open('gpurun_out/r02v_bench_${N}gpu.json')); print({k:d[k] for k in ('value','ms_per_step','frames','parity_check','e2e','throughput_mode')})
PY
