#!/bin/bash
# multi-GPU bench under torchrun: N = $1 ranks (weak-scaling headline line + strong-scaling configs[4] leg)
N=${1:-2}
O=gpurun_out/r02s; mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > $O/bench_${N}gpu.json 2> $O/bench_${N}gpu.err; echo "exit $?"; tail -5 $O/bench_${N}gpu.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $O/bench_ref_${N}gpu.json 2> $O/bench_ref_${N}gpu.err; echo "ref exit $?"
python - <<PY
import json
d=json.loads(open('$O/bench_${N}gpu.json').read().strip().splitlines()[-1])
for k in ('value','n_gpus','ms_per_step','e2e','configs4','clocks'): print(k, json.dumps(d.get(k))[:900])
PY
tail -c 300 $O/bench_ref_${N}gpu.json
