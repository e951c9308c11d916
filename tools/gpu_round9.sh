#!/bin/bash
set -u
mkdir -p gpurun_out
export B200PT_QUIET=1
bash tools/ab2.sh
