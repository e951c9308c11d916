#!/bin/bash
# round-2 GPU pass G2 (2 GPUs): bench at N=2 launched like the driver does
export B200PT_QUIET=1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/r02g_bench_2gpu.json 2> gpurun_out/r02g_bench_2gpu.err; echo "bench rc=$?"; tail -3 gpurun_out/r02g_bench_2gpu.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02g_bench_2gpu.json')); print({k:d[k] for k in ('value','ms_per_step','frames','parity_check','e2e','throughput_mode')})
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 | cut -c1-300
