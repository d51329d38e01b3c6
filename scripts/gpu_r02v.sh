#!/bin/bash
# r02v: A/B of the psi sweep (P0 two sweeps / P1 one sweep) and the pre-beam threshold (T0 min rule / T1 tight rule)
mkdir -p gpurun_out/r02v
L=llm_guided_asr_b200/csrc
cp $L/liblgasr.so /tmp/liblgasr_keep.so
for v in P0T0 P1T0 P0T1 P1T1; do
  cp $L/variants/liblgasr_$v.so $L/liblgasr.so
  echo "== $v" >> gpurun_out/r02v/step_probe.txt
  timeout 300 python scripts/step_probe.py --B 64 512 >> gpurun_out/r02v/step_probe.txt 2>&1
done
cp /tmp/liblgasr_keep.so $L/liblgasr.so
timeout 600 python -m pytest tests/test_gpu_fused_search.py tests/test_gpu_edge_cases.py tests/test_gpu_baseline_sizes.py -q -x 2>&1 | tail -5
cut -c1-175 gpurun_out/r02v/step_probe.txt
