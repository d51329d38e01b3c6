#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export LGASR_V2=1
for n in 128 64; do
  LGASR_NTH=$n timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02e_pytest_$n.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02e_pytest_$n.log
  tail -4 gpurun_out/r02e_pytest_$n.log
done
for n in 256 128 64; do
  echo "NTH=$n" >> gpurun_out/r02e_step_probe.txt
  LGASR_NTH=$n timeout 300 python scripts/step_probe.py --B 64 128 512 >> gpurun_out/r02e_step_probe.txt 2>&1
done
cat gpurun_out/r02e_step_probe.txt
LGASR_NTH=128 timeout 120 python scripts/trace_step2.py 64 > gpurun_out/r02e_trace64_128.txt 2>&1; cat gpurun_out/r02e_trace64_128.txt
LGASR_NTH=64 timeout 120 python scripts/trace_step2.py 64 > gpurun_out/r02e_trace64_64.txt 2>&1; cat gpurun_out/r02e_trace64_64.txt
LGASR_NTH=64 timeout 300 python scripts/step_probe.py --B 512 --mode step --reps 2 > gpurun_out/r02e_plain.log 2>&1 && \
LGASR_NTH=64 timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_step2 -s 14 -c 1 -o gpurun_out/r02e_step512_64 \
  python scripts/step_probe.py --B 512 --mode step --reps 2 > gpurun_out/r02e_ncu.log 2>&1
