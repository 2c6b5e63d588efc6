#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "varchar or string or config3 or dictionary or golden_strings" > gpurun_out/r1p_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r1p_tests.log
tail -12 gpurun_out/r1p_tests.log
SWEEP_TYPE=varchar timeout 400 python tools/probe_sweep.py 2e8 3 8e4,1e7,5e7,1e8 > gpurun_out/r1p_varchar.txt 2>&1
cat gpurun_out/r1p_varchar.txt
