#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_device_cache.py tests/test_gpu_reference_driving.py -m gpu -q > gpurun_out/r02o_pytest.log 2>&1; tail -15 gpurun_out/r02o_pytest.log | cut -c1-300
timeout 1200 python bench.py --steps 5 --warmup 3 --no-configs4 > gpurun_out/r02o_bench.json 2> gpurun_out/r02o_bench.err; echo "bench exit $?"; tail -5 gpurun_out/r02o_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02o_bench.json').read().strip().splitlines()[-1])
for k in ('decode_real',): print(k, json.dumps(d.get(k))[:1200])
PY
