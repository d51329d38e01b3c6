#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export LGASR_V2=1
for n in 256 128; do
LGASR_NTH=$n timeout 120 python scripts/trace_step2.py 64 > gpurun_out/r02f_trace64_$n.txt 2>&1; cat gpurun_out/r02f_trace64_$n.txt
done
LGASR_NTH=256 timeout 120 python scripts/trace_step2.py 8 > gpurun_out/r02f_trace8_256.txt 2>&1; cat gpurun_out/r02f_trace8_256.txt
