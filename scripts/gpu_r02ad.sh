#!/bin/bash
O=gpurun_out/r02ad; mkdir -p $O
L=llm_guided_asr_b200/csrc
cp $L/liblgasr.so /tmp/keep.so
for v in g10 g5; do
  cp $L/variants/liblgasr_$v.so $L/liblgasr.so
  echo "== $v (dense forced)" >> $O/probe.txt
  LGASR_V2=1 LGASR_DENSE=1 timeout 300 python scripts/step_probe.py --B 128 512 --K 1 >> $O/probe.txt 2>&1
done
cp /tmp/keep.so $L/liblgasr.so
cut -c1-175 $O/probe.txt
