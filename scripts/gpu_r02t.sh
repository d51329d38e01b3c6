#!/bin/bash
# r02t: full GPU suite + bench after the sharded-driver change
mkdir -p gpurun_out/r02t
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02t/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02t/pytest.log
tail -5 gpurun_out/r02t/pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02t/bench.json 2> gpurun_out/r02t/bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r02t/bench.json
