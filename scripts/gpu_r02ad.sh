#!/bin/bash
O=gpurun_out/r02ag; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log; tail -3 $O/pytest.log
LGASR_NO_CLUSTER=1 timeout 900 python -m pytest tests/test_gpu_fused_search.py tests/test_gpu_edge_cases.py tests/test_gpu_device_cache.py -q 2>&1 | tail -2
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02ag/bench.json').read().strip().splitlines()[-1])
r=d['roofline']; print('step',r['avg_launch_us'],r['frac'],'replay',r['replay']['avg_launch_us'],r['replay']['frac'])
print('b512',d['roofline_b512']['avg_launch_us'],d['roofline_b512']['frac'])
print({k:(round(v['ms_per_position'],3),v['split_ms_per_position']) for k,v in d['decode_real'].items() if isinstance(v,dict)})
print(d['parity'], d['value'])
PY
