#!/bin/bash
# round-2 GPU pass S: pooled triangle phase (PT_TRI_POOL) against the product walker
export B200PT_QUIET=1
L=owl_path_tracer_b200/lib
timeout 900 python tools/ab.py --spp 32 --scenes dragon,car --timing --check base=$L/libb200pt.so npool=$L/variants/libb200pt_npool.so > gpurun_out/r02s_ab7.jsonl 2> gpurun_out/r02s_ab7.err; echo "ab rc=$?"
cut -c1-300 gpurun_out/r02s_ab7.jsonl
tail -3 gpurun_out/r02s_ab7.err
