#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_window_online.py tests/test_gpu_stream.py -m gpu -q > gpurun_out/r02r_pytest.log 2>&1; tail -30 gpurun_out/r02r_pytest.log | cut -c1-300
