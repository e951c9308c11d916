#!/bin/bash
# round-2 GPU pass N: reinsertion passes in the builder -- gpu tests, then A/B of 0 / 4 (product) / 8 passes
export B200PT_QUIET=1
L=owl_path_tracer_b200/lib
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 900 python tools/ab.py --spp 32 --scenes dragon,car --timing --check base=$L/libb200pt.so r0=$L/variants/libb200pt_r0.so r8=$L/variants/libb200pt_r8.so > gpurun_out/r02n_ab.jsonl 2> gpurun_out/r02n_ab.err; echo "ab rc=$?"
cut -c1-600 gpurun_out/r02n_ab.jsonl
tail -3 gpurun_out/r02n_ab.err
