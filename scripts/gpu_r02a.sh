#!/bin/bash
# round-2 first GPU call: baseline state of the round-1 build (tests, stepping/replay launch times, ncu source-level capture)
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02a_pytest.log
tail -3 gpurun_out/r02a_pytest.log
timeout 300 python scripts/step_probe.py --B 64 128 512 > gpurun_out/r02a_step_probe.txt 2>&1
cat gpurun_out/r02a_step_probe.txt
timeout 300 python scripts/step_probe.py --B 64 --mode step --reps 2 > gpurun_out/r02a_plain64.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_step -s 14 -c 2 -o gpurun_out/r02a_step64 \
  python scripts/step_probe.py --B 64 --mode step --reps 2 > gpurun_out/r02a_ncu64.log 2>&1
timeout 300 python scripts/step_probe.py --B 512 --mode step --reps 2 > gpurun_out/r02a_plain512.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_step -s 14 -c 1 -o gpurun_out/r02a_step512 \
  python scripts/step_probe.py --B 512 --mode step --reps 2 > gpurun_out/r02a_ncu512.log 2>&1
ls -la gpurun_out
