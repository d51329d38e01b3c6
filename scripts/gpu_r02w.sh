#!/bin/bash
# r02w: evidence run of the round-2 build: GPU suite, bench (ours + reference arm), then the ncu launch list of the bench
O=gpurun_out/r02w; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log; tail -3 $O/pytest.log
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu --no-configs4 > $O/bench_short.json 2> $O/bench_short.err; rc=$?; echo "short rc=$rc"
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-configs4 > $O/ncu_bench.log 2>&1; echo "ncu rc=$?"
fi
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02w/bench.json').read().strip().splitlines()[-1])
r=d['roofline']
print('value',d['value'],'e2e',d['e2e']['value'],'launches',d['gpu_launches'])
print('step',r['avg_launch_us'],r['frac'],'replay',r['replay']['avg_launch_us'],r['replay']['frac'])
print('b512',d['roofline_b512']['avg_launch_us'],d['roofline_b512']['frac'],d['roofline_b512']['replay'])
print('c4',d['configs4']['value'] if d['configs4'] else None,'cpu',d['cpu_baseline']['value'],'refgpu',d['reference_gpu'].get('value'),'parity',d['parity'])
print('clocks',d['clocks'])
PY
tail -c 600 $O/bench_ref.json
