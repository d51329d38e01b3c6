#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export LGASR_V3=1
run() { echo "== $*"; timeout 120 python scripts/step_probe.py "$@" 2>&1 | tail -2 | cut -c1-300; }
run --B 128 --L 8 --mode step
run --B 256 --L 8 --mode step
run --B 512 --L 8 --mode step
run --B 1024 --L 6 --mode step
