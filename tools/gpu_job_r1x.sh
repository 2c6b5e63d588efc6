#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "config1 or filter_sizes or flat_columns or ragged or types_flat or golden or known" > gpurun_out/r1x_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r1x_tests.log
tail -4 gpurun_out/r1x_tests.log
{
timeout 300 python tools/probe_sweep.py 1e9 5 1e3,1e4,4e4,8e4,1e7
SWEEP_TYPE=int32 timeout 300 python tools/probe_sweep.py 1e9 5 8e4
} > gpurun_out/r1x_sweep.txt 2>&1
cat gpurun_out/r1x_sweep.txt
