#!/bin/bash
# Short closing check of the shipped library: variant tests, smoke (incl. the variant), small-filter sweep
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_acero.py -m gpu -q > gpurun_out/r1z5_acero_tests.log 2>&1
echo "acero tests rc=$?" >> gpurun_out/r1z5_acero_tests.log
tail -3 gpurun_out/r1z5_acero_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r1z5_smoke.log 2>&1
tail -3 gpurun_out/r1z5_smoke.log
timeout 300 python tools/acero_sweep.py 1e9 5 2e4,8e4,1e7 1 > gpurun_out/r1z5_acero_sweep.txt 2>&1
cat gpurun_out/r1z5_acero_sweep.txt
