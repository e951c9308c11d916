#!/bin/bash
export B200PT_QUIET=1
V=owl_path_tracer_b200/lib/variants
timeout 600 python tools/ab.py --spp 32 --timing base=owl_path_tracer_b200/lib/libb200pt.so trigd=$V/libb200pt_trigd.so base2=owl_path_tracer_b200/lib/libb200pt.so > gpurun_out/r02f_ab.jsonl 2>&1; cut -c1-260 gpurun_out/r02f_ab.jsonl
(timeout 900 python tools/trig_probe.py owl_path_tracer_b200/lib/libb200pt.so; timeout 900 python tools/trig_probe.py $V/libb200pt_trigd.so) > gpurun_out/r02f_trig_probe.jsonl 2>&1; cat gpurun_out/r02f_trig_probe.jsonl
