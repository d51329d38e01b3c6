#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r02m_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02m_pytest.log
tail -40 gpurun_out/r02m_pytest.log | cut -c1-300
