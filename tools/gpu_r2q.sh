#!/bin/bash
# round-2 GPU pass Q: out-of-L2 mesh (7.2 M triangles) after the builder change -- the plain command, then one ncu --set full capture of k_extend
export B200PT_QUIET=1
B="python bench.py --detail 7200000 --pipeline 1 --streams-mode 0 --no-cpu-baseline --steps 1 --warmup 0 --spp 16 --no-parity-check"
timeout 900 $B > gpurun_out/r02q_detail_spp16.json 2> gpurun_out/r02q_detail_spp16.err; echo "rc=$?"; cut -c1-600 gpurun_out/r02q_detail_spp16.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_extend -s 4 -c 3 -f -o gpurun_out/extend_detail $B > gpurun_out/ncu_extend_detail.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
