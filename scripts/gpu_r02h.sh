#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_reference_driving.py -m gpu -q > gpurun_out/r02h_pytest.log 2>&1; tail -5 gpurun_out/r02h_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02h_bench.json 2> gpurun_out/r02h_bench.err; echo "bench exit $?"; tail -5 gpurun_out/r02h_bench.err; cat gpurun_out/r02h_bench.json
