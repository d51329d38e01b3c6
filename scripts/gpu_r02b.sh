#!/bin/bash
# round-2: first run of k_step2 (parity tests + launch times, both kernels)
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02b_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02b_pytest.log
tail -25 gpurun_out/r02b_pytest.log
timeout 300 python scripts/step_probe.py --B 64 128 512 > gpurun_out/r02b_step_probe.txt 2>&1
cat gpurun_out/r02b_step_probe.txt
LGASR_V1=1 timeout 300 python scripts/step_probe.py --B 64 512 > gpurun_out/r02b_step_probe_v1.txt 2>&1
cat gpurun_out/r02b_step_probe_v1.txt
timeout 300 python scripts/step_probe.py --B 64 --K 2 > gpurun_out/r02b_step_probe_k2.txt 2>&1
timeout 300 python scripts/step_probe.py --B 32 128 --K 2 --shape owsm >> gpurun_out/r02b_step_probe_k2.txt 2>&1
cat gpurun_out/r02b_step_probe_k2.txt
