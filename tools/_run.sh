#!/bin/bash
T=${1:-t01}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -x > gpurun_out/${T}_alone.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/${T}_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_all.log 2>&1; echo "all rc=$?"; tail -4 gpurun_out/${T}_all.log
