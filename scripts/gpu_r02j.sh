#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export LGASR_V3=1
run() { echo "== $*"; timeout 120 python scripts/step_probe.py "$@" 2>&1 | tail -2 | cut -c1-300; }
run --B 8 --L 12 --mode step
run --B 8 --L 12 --mode replay
run --B 16 --L 6 --mode step
run --B 64 --L 4 --mode step
run --B 64 --L 4 --mode replay
run --B 64 --L 12 --mode step
run --B 64 --L 12 --mode replay
