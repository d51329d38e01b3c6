#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export LGASR_V2=1
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02d_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02d_pytest.log
tail -8 gpurun_out/r02d_pytest.log
timeout 300 python scripts/step_probe.py --B 64 128 512 > gpurun_out/r02d_step_probe.txt 2>&1; cat gpurun_out/r02d_step_probe.txt
timeout 120 python scripts/trace_step2.py 64 > gpurun_out/r02d_trace64.txt 2>&1; cat gpurun_out/r02d_trace64.txt
timeout 300 python scripts/step_probe.py --B 64 --mode step --reps 2 > gpurun_out/r02d_plain64.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_step2 -s 14 -c 2 -o gpurun_out/r02d_step64 \
  python scripts/step_probe.py --B 64 --mode step --reps 2 > gpurun_out/r02d_ncu64.log 2>&1
ls gpurun_out | tail -3
