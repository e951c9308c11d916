#!/bin/bash
# round-2 GPU pass S: pooled triangle phase (PT_TRI_POOL) against the product walker
export B200PT_QUIET=1
L=owl_path_tracer_b200/lib
timeout 900 python tools/ab.py --spp 32 --scenes dragon,car --timing --check base=$L/libb200pt.so rf3=$L/variants/libb200pt_rf3.so rf10=$L/variants/libb200pt_rf10.so sl6=$L/variants/libb200pt_sl6.so sl1=$L/variants/libb200pt_sl1.so > gpurun_out/r02s_ab5.jsonl 2> gpurun_out/r02s_ab5.err; echo "ab rc=$?"
cut -c1-300 gpurun_out/r02s_ab5.jsonl
tail -3 gpurun_out/r02s_ab5.err
