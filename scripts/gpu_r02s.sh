#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02s_bench2.json 2> gpurun_out/r02s_bench2.err; echo "exit $?"; tail -5 gpurun_out/r02s_bench2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02s_bench2.json').read().strip().splitlines()[-1])
for k in ('value','n_gpus','ms_per_step','e2e','configs4','clocks'): print(k, json.dumps(d.get(k))[:700])
PY
