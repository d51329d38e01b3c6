#!/bin/bash
O=gpurun_out/r02aj; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_device_cache.py tests/test_gpu_reference_driving.py tests/test_gpu_llm_handoff.py tests/test_gpu_stream.py tests/test_reference_surface.py -m gpu -x -q > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log; tail -12 $O/pytest.log
timeout 900 python bench.py --steps 3 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02aj/bench.json').read().strip().splitlines()[-1])
print({k:(round(v['ms_per_position'],3),{a:round(b,3) for a,b in v['split_ms_per_position'].items()}) for k,v in d['decode_real'].items() if isinstance(v,dict)})
PY
