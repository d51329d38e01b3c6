#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r02g_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02g_pytest.log
tail -30 gpurun_out/r02g_pytest.log
LGASR_V2=1 LGASR_NTH=256 timeout 1200 python -m pytest tests/test_gpu_baseline_sizes.py tests/test_gpu_reference_driving.py tests/test_gpu_edge_cases.py -m gpu -q > gpurun_out/r02g_pytest_v2.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02g_pytest_v2.log
tail -12 gpurun_out/r02g_pytest_v2.log
nproc
timeout 600 python bench.py --impl reference_gpu --steps 5 > gpurun_out/r02g_refgpu.json 2> gpurun_out/r02g_refgpu.err; cut -c1-600 gpurun_out/r02g_refgpu.json; tail -3 gpurun_out/r02g_refgpu.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02g_ref.json 2> gpurun_out/r02g_ref.err; cut -c1-400 gpurun_out/r02g_ref.json; tail -3 gpurun_out/r02g_ref.err
