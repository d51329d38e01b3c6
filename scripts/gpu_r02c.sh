#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02c_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02c_pytest.log
tail -8 gpurun_out/r02c_pytest.log
timeout 120 python scripts/trace_step2.py 64 > gpurun_out/r02c_trace64.txt 2>&1; cat gpurun_out/r02c_trace64.txt
TRACE_REPLAY=1 timeout 120 python scripts/trace_step2.py 64 > gpurun_out/r02c_trace64_replay.txt 2>&1; cat gpurun_out/r02c_trace64_replay.txt
timeout 120 python scripts/trace_step2.py 512 > gpurun_out/r02c_trace512.txt 2>&1; cat gpurun_out/r02c_trace512.txt
timeout 300 python scripts/step_probe.py --B 64 --mode step --reps 2 > gpurun_out/r02c_plain64.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_step2 -s 14 -c 2 -o gpurun_out/r02c_step64 \
  python scripts/step_probe.py --B 64 --mode step --reps 2 > gpurun_out/r02c_ncu64.log 2>&1
ls -la gpurun_out | tail -5
