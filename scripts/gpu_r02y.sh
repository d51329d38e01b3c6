#!/bin/bash
# r02y: ncu --set full capture of the stepping k_step instance at B = 512, after the same command ran clean
O=gpurun_out/r02y; mkdir -p $O
timeout 300 python scripts/step_probe.py --B 512 --mode step --reps 2 > $O/plain512.log 2>&1; rc=$?; cat $O/plain512.log | cut -c1-200
if [ $rc -eq 0 ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 14 -c 1 -o $O/step512 \
    python scripts/step_probe.py --B 512 --mode step --reps 2 > $O/ncu512.log 2>&1; echo "ncu rc=$?"
fi
ls -la $O
