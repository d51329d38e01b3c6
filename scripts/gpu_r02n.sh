#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_device_cache.py tests/test_gpu_fused_search.py -m gpu -q -x > gpurun_out/r02n_pytest.log 2>&1; tail -5 gpurun_out/r02n_pytest.log
timeout 1200 python bench.py --steps 5 --warmup 3 > gpurun_out/r02n_bench.json 2> gpurun_out/r02n_bench.err; echo "bench exit $?"; tail -5 gpurun_out/r02n_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02n_bench.json').read().strip().splitlines()[-1])
for k in ('value','e2e','configs4','decode_real','clocks'): print(k, json.dumps(d.get(k))[:1200])
PY
