#!/bin/bash
# ncu evidence: launch list of the bench command + full captures of k_extend / k_shade (dragon) and k_shade (car).
# Every ncu pass runs only after the same command exited 0 without ncu.  Numbers printed under ncu are never bench values.
set -u
export B200PT_QUIET=1
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --pipeline 1 --streams-mode 0 --no-cpu-baseline --no-parity-check"
timeout 600 $B --spp 16 > gpurun_out/bench_spp16.log 2>&1 || { echo "plain bench failed"; tail -5 gpurun_out/bench_spp16.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches.csv $B --spp 8 > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_extend -s 4 -c 3 -f -o gpurun_out/extend $B --spp 16 > gpurun_out/ncu_extend.log 2>&1
echo "extend capture rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 2 -c 2 -f -o gpurun_out/shade $B --spp 16 > gpurun_out/ncu_shade.log 2>&1
echo "shade capture rc=$?"
ls -la gpurun_out | head -30
