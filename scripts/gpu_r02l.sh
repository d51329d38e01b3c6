#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_llm_handoff.py tests/test_gpu_reference_driving.py -m gpu -q > gpurun_out/r02l_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02l_pytest.log
tail -60 gpurun_out/r02l_pytest.log | cut -c1-400
