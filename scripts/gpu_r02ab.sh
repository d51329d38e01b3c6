#!/bin/bash
O=gpurun_out/r02ab; mkdir -p $O
timeout 600 python scripts/decoder_profile.py > $O/decoder_profile.txt 2>&1; tail -32 $O/decoder_profile.txt | cut -c1-200
timeout 900 python bench.py --steps 3 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -3 $O/bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02ab/bench.json').read().strip().splitlines()[-1])
for r in d['roofline_other_shapes']: print(json.dumps(r)[:700])
PY
