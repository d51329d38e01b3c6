#!/bin/bash
O=gpurun_out/r02am; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log; tail -4 $O/pytest.log
LGASR_V2=1 timeout 900 python -m pytest tests/test_gpu_fused_search.py tests/test_gpu_edge_cases.py tests/test_gpu_baseline_sizes.py -q -x 2>&1 | tail -2
timeout 300 python scripts/step_probe.py --B 64 128 512 > $O/probe.txt 2>&1
timeout 300 python scripts/step_probe.py --B 64 512 --K 2 >> $O/probe.txt 2>&1
cut -c1-150 $O/probe.txt
LGASR_TRACE=1 timeout 300 python scripts/trace_step.py 2>&1 | grep "select CTAs\|select_done\|ticket " | head -4
