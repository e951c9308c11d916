#!/bin/bash
# round-2 GPU pass A: the no-allocation entry point alone, the whole GPU suite, smoke, a default bench run
export B200PT_QUIET=1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -8
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; echo "bench rc=$?"; cat gpurun_out/r02a_bench.json; tail -3 gpurun_out/r02a_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02a_ref.json 2> gpurun_out/r02a_ref.err; echo "ref rc=$?"; cat gpurun_out/r02a_ref.json
