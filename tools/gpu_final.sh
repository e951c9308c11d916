#!/bin/bash
# round-end evidence: gpu tests, full default bench + reference arm, ncu launch list + captures, scene sweep
set -u
mkdir -p gpurun_out
B200PT_QUIET=1 timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 1500 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench_full.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference rc=$?"; tail -c 400 gpurun_out/bench_reference.json
bash tools/gpu_ncu.sh > gpurun_out/ncu_run.log 2>&1; tail -3 gpurun_out/ncu_run.log
timeout 1200 python tools/scene_sweep.py > gpurun_out/scene_sweep.log 2>&1; echo "sweep rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
