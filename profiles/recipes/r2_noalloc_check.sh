#!/bin/bash
# the no-alloc discretiser: its test alone in a fresh process, then the whole GPU suite (it runs last)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/na_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/na_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/na_suite.log 2>&1; echo "suite rc=$?" >> gpurun_out/na_suite.log
tail -n 3 gpurun_out/na_alone.log gpurun_out/na_suite.log
