#!/bin/bash
# r02x: ncu --set full capture of the stepping k_step instance at B = 64 (configs[1] shape), after the same command ran clean
O=gpurun_out/r02x; mkdir -p $O
timeout 300 python scripts/step_probe.py --B 64 --mode step --reps 2 > $O/plain64.log 2>&1; rc=$?; cat $O/plain64.log | cut -c1-200
if [ $rc -eq 0 ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 14 -c 2 -o $O/step64 \
    python scripts/step_probe.py --B 64 --mode step --reps 2 > $O/ncu64.log 2>&1; echo "ncu rc=$?"
fi
ls -la $O
