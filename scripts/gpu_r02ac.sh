#!/bin/bash
O=gpurun_out/r02ac; mkdir -p $O
( time python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" ) > $O/smoke.log 2>&1; tail -6 $O/smoke.log
( time python bench.py > $O/bench.json 2> $O/bench.err ) 2> $O/bench_time.log; echo "bench rc=$?"; cat $O/bench_time.log | tail -4
( time python bench.py --impl reference > $O/bench_ref.json 2> $O/bench_ref.err ) 2> $O/ref_time.log; echo "ref rc=$?"; tail -4 $O/ref_time.log
