#!/bin/bash
# last check of the round: GPU suite, smoke, a short bench line
set -u
mkdir -p gpurun_out
export B200PT_QUIET=1
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -v "^ - \|^Parsing\|^Image" gpurun_out/pytest_gpu.log | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --steps 1 --warmup 3 --spp 128 --no-cpu-baseline --streams-mode 0 2>/dev/null | cut -c1-400
