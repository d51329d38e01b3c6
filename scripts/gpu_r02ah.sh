#!/bin/bash
# r02ah: the no-alloc entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last)
O=gpurun_out/r02ah; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest_alone.log 2>&1; echo "alone rc=$?" >> $O/pytest_alone.log; tail -15 $O/pytest_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_all.log 2>&1; echo "all rc=$?" >> $O/pytest_all.log; tail -5 $O/pytest_all.log
