#!/bin/bash
# ncu evidence: launch list of the bench command + full captures of k_extend / k_shade.
set -u
mkdir -p gpurun_out
timeout 600 python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline > gpurun_out/bench_spp64.log 2>&1 || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --spp 16 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_extend -s 4 -c 3 -f -o gpurun_out/extend \
    python bench.py --steps 1 --warmup 0 --spp 16 --no-cpu-baseline > gpurun_out/ncu_extend.log 2>&1
echo "extend capture rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 2 -c 2 -f -o gpurun_out/shade \
    python bench.py --steps 1 --warmup 0 --spp 16 --no-cpu-baseline > gpurun_out/ncu_shade.log 2>&1
echo "shade capture rc=$?"
ls -la gpurun_out
