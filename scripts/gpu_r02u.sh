#!/bin/bash
# r02u: v1 k_step with one-sweep psi + tight pre-beam threshold: parity suite, then launch times
mkdir -p gpurun_out/r02u
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02u/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02u/pytest.log
tail -4 gpurun_out/r02u/pytest.log
timeout 300 python scripts/step_probe.py --B 64 128 512 > gpurun_out/r02u/step_probe.txt 2>&1
timeout 300 python scripts/step_probe.py --B 64 512 --K 2 >> gpurun_out/r02u/step_probe.txt 2>&1
timeout 300 python scripts/step_probe.py --B 32 128 --K 2 --shape owsm --L 6 >> gpurun_out/r02u/step_probe.txt 2>&1
cat gpurun_out/r02u/step_probe.txt | cut -c1-200
