#!/bin/bash
# round-2 GPU pass C: float32 deterministic trig (goldens regenerated) — whole GPU suite, bench with the frame pipeline
export B200PT_QUIET=1
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
timeout 900 python bench.py --steps 6 --warmup 3 > gpurun_out/r02c_bench.json 2> gpurun_out/r02c_bench.err; echo "bench rc=$?"; cat gpurun_out/r02c_bench.json; tail -3 gpurun_out/r02c_bench.err
timeout 600 python bench.py --steps 6 --warmup 3 --pipeline 3 --start-fraction 0.6 --no-cpu-baseline --streams-mode 0 > gpurun_out/r02c_bench_p3.json 2> gpurun_out/r02c_bench_p3.err; echo "bench rc=$?"; cut -c1-900 gpurun_out/r02c_bench_p3.json
timeout 600 python tools/ab.py --spp 32 --timing base=owl_path_tracer_b200/lib/libb200pt.so > gpurun_out/r02c_ab.jsonl 2>&1; cat gpurun_out/r02c_ab.jsonl | cut -c1-330
