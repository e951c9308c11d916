#!/bin/bash
export B200PT_QUIET=1
V=owl_path_tracer_b200/lib/variants
timeout 900 python tools/ab.py --spp 32 --timing --check base=owl_path_tracer_b200/lib/libb200pt.so shade3=$V/libb200pt_shade3.so shade5=$V/libb200pt_shade5.so shade6=$V/libb200pt_shade6.so libm=$V/libb200pt_libm.so > gpurun_out/r02d_ab.jsonl 2>&1; cut -c1-330 gpurun_out/r02d_ab.jsonl
