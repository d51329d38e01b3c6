#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export LGASR_V3=1
timeout 300 python scripts/step_probe.py --B 8 --L 6 > gpurun_out/r02i_probe_small.txt 2>&1; cat gpurun_out/r02i_probe_small.txt
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02i_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02i_pytest.log
tail -30 gpurun_out/r02i_pytest.log
timeout 300 python scripts/step_probe.py --B 64 128 512 > gpurun_out/r02i_step_probe.txt 2>&1; cat gpurun_out/r02i_step_probe.txt
