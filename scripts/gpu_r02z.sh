#!/bin/bash
# r02z: one-warp selection tail: parity suite, launch times, phase trace
O=gpurun_out/r02z; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log; tail -3 $O/pytest.log
LGASR_NO_CLUSTER=1 timeout 900 python -m pytest tests/test_gpu_fused_search.py tests/test_gpu_edge_cases.py -q -x 2>&1 | tail -2
timeout 300 python scripts/step_probe.py --B 64 128 512 > $O/step_probe.txt 2>&1
timeout 300 python scripts/step_probe.py --B 64 512 --K 2 >> $O/step_probe.txt 2>&1
cut -c1-175 $O/step_probe.txt
LGASR_TRACE=1 timeout 300 python scripts/trace_step.py > $O/trace.txt 2>&1; tail -30 $O/trace.txt
