#!/bin/bash
O=gpurun_out/r02ai; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log; tail -4 $O/pytest.log
timeout 900 python bench.py --impl reference_gpu --steps 5 > $O/refgpu.json 2> $O/refgpu.err; echo "rc=$?"; tail -3 $O/refgpu.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02ai/refgpu.json').read().strip().splitlines()[-1])
print(d['ms_per_position'], json.dumps(d['scorer_only_adoption'],indent=1))
PY
