#!/bin/bash
# r02aa: ncu --set full (+ source) of the dense k_step2 instance, configs[4] per-GPU batch (B = 512, two matrices), stepping
O=gpurun_out/r02aa; mkdir -p $O
timeout 300 python scripts/step_probe.py --B 512 --K 2 --mode step --reps 2 > $O/plain.log 2>&1; rc=$?; cut -c1-200 $O/plain.log
if [ $rc -eq 0 ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step2 -s 14 -c 1 -o $O/step2_512k2 \
    python scripts/step_probe.py --B 512 --K 2 --mode step --reps 2 > $O/ncu.log 2>&1; echo "ncu rc=$?"
fi
ls -la $O
