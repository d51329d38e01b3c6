#!/bin/bash
O=gpurun_out/r02ae; mkdir -p $O
for dl in 1 2; do
  echo "== LGASR_DENSE=$dl" >> $O/probe.txt
  LGASR_V2=1 LGASR_DENSE=$dl timeout 300 python scripts/step_probe.py --B 128 512 --K 1 >> $O/probe.txt 2>&1
  LGASR_V2=1 LGASR_DENSE=$dl timeout 300 python scripts/step_probe.py --B 512 --K 2 >> $O/probe.txt 2>&1
done
cut -c1-175 $O/probe.txt
LGASR_V2=1 LGASR_DENSE=2 timeout 900 python -m pytest tests/test_gpu_fused_search.py tests/test_gpu_edge_cases.py tests/test_gpu_baseline_sizes.py -q 2>&1 | tail -3
