#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stream.py -m gpu -q -x > gpurun_out/r02p_pytest.log 2>&1; tail -25 gpurun_out/r02p_pytest.log | cut -c1-300
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r02p_pytest_all.log 2>&1; tail -6 gpurun_out/r02p_pytest_all.log | cut -c1-300
