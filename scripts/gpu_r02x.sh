#!/bin/bash
# r02x: OWSM / two-matrix shapes: k_step (default) against k_step2 (LGASR_V2=1), both launch modes; configs[4] phase probe
O=gpurun_out/r02x; mkdir -p $O
for v in 0 1; do
  echo "== LGASR_V2=$v" >> $O/shape_probe.txt
  LGASR_V2=$v timeout 300 python scripts/step_probe.py --B 32 128 --K 2 --shape owsm --L 6 >> $O/shape_probe.txt 2>&1
  LGASR_V2=$v timeout 300 python scripts/step_probe.py --B 32 128 --K 1 --shape owsm --L 6 >> $O/shape_probe.txt 2>&1
  LGASR_V2=$v timeout 300 python scripts/step_probe.py --B 64 512 --K 2 >> $O/shape_probe.txt 2>&1
done
cut -c1-175 $O/shape_probe.txt
timeout 300 python scripts/configs4_probe.py > $O/configs4_probe.txt 2>&1; tail -6 $O/configs4_probe.txt
