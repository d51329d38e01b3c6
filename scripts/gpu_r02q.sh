#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r02q_pytest.log 2>&1; tail -30 gpurun_out/r02q_pytest.log | cut -c1-300
