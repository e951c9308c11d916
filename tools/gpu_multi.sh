#!/bin/bash
# N-GPU bench line exactly as the driver launches it (N from the first argument)
set -u
N=${1:-2}
mkdir -p gpurun_out
export B200PT_QUIET=1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err
echo "bench $N rc=$?"; tail -c 1500 gpurun_out/bench_${N}gpu.json; tail -5 gpurun_out/bench_${N}gpu.err
