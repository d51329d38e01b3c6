#!/bin/bash
# r02y3: dense k_step2 instance: whole GPU suite with it forced on every launch, then the default routing (suite + probes)
O=gpurun_out/r02y3; mkdir -p $O
LGASR_V2=1 LGASR_DENSE=1 timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_dense.log 2>&1; echo "rc=$?" >> $O/pytest_dense.log; tail -4 $O/pytest_dense.log
LGASR_V2=1 timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_v2.log 2>&1; echo "rc=$?" >> $O/pytest_v2.log; tail -3 $O/pytest_v2.log
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log; tail -3 $O/pytest.log
timeout 300 python scripts/step_probe.py --B 64 512 --K 2 > $O/probe.txt 2>&1
timeout 300 python scripts/step_probe.py --B 64 512 --K 1 >> $O/probe.txt 2>&1
timeout 300 python scripts/step_probe.py --B 32 128 --K 2 --shape owsm --L 6 >> $O/probe.txt 2>&1
cut -c1-175 $O/probe.txt
