#!/bin/bash
# r02ak: fresh ncu --set full capture of the stepping k_step instance (B = 64, configs[1] shape) of the final build
O=gpurun_out/r02ak; mkdir -p $O
timeout 300 python scripts/step_probe.py --B 64 --mode step --reps 2 > $O/plain64.log 2>&1; rc=$?; cut -c1-200 $O/plain64.log
if [ $rc -eq 0 ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 14 -c 2 -o $O/step64 \
    python scripts/step_probe.py --B 64 --mode step --reps 2 > $O/ncu64.log 2>&1; echo "ncu rc=$?"
fi
ls -la $O
