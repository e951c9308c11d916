#!/bin/bash
set -u
mkdir -p gpurun_out
export B200PT_QUIET=1
bash tools/ab2.sh | grep "^==\|1080\|640\|rror" | cut -c1-300
