#!/bin/bash
# round-2 GPU pass H: the out-of-L2 mesh (7.2 M triangles) — a short bench run, then one ncu --set full capture of k_extend
export B200PT_QUIET=1
B="python bench.py --detail 7200000 --pipeline 1 --streams-mode 0 --no-cpu-baseline"
timeout 900 $B --steps 2 --warmup 1 --spp 64 > gpurun_out/r02h_detail_spp64.json 2> gpurun_out/r02h_detail_spp64.err; echo "rc=$?"; cut -c1-1800 gpurun_out/r02h_detail_spp64.json; tail -2 gpurun_out/r02h_detail_spp64.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_extend -s 4 -c 3 -f -o gpurun_out/extend_detail $B --steps 1 --warmup 0 --spp 16 --no-parity-check > gpurun_out/ncu_extend_detail.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
