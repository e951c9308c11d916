#!/bin/bash
# round-2 GPU pass S: pooled triangle phase (PT_TRI_POOL) against the product walker
export B200PT_QUIET=1
L=owl_path_tracer_b200/lib
timeout 900 python tools/ab.py --spp 32 --scenes dragon,car --timing --check pool6=$L/variants/libb200pt_pool6.so c10=$L/variants/libb200pt_c10.so c06=$L/variants/libb200pt_c06.so c05=$L/variants/libb200pt_c05.so c04=$L/variants/libb200pt_c04.so > gpurun_out/r02s_ab4.jsonl 2> gpurun_out/r02s_ab4.err; echo "ab rc=$?"
cut -c1-300 gpurun_out/r02s_ab4.jsonl
tail -3 gpurun_out/r02s_ab4.err
