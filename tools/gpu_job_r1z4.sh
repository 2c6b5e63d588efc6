#!/bin/bash
# Round-1 closing run on the shipped library: variant tests + sweep, smoke (now incl. the variant), default bench
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_acero.py -m gpu -q > gpurun_out/r1z4_acero_tests.log 2>&1
echo "acero tests rc=$?" >> gpurun_out/r1z4_acero_tests.log
tail -5 gpurun_out/r1z4_acero_tests.log
timeout 600 python tools/acero_sweep.py 1e9 5 2e4,8e4,1e7,1e8 1,16 > gpurun_out/r1z4_acero_sweep.txt 2>&1
cat gpurun_out/r1z4_acero_sweep.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r1z4_smoke.log 2>&1
tail -3 gpurun_out/r1z4_smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r1z4_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r1z4_tests.log
tail -3 gpurun_out/r1z4_tests.log
timeout 600 python bench.py > gpurun_out/r1z4_bench.json 2> gpurun_out/r1z4_bench.err
cut -c1-300 gpurun_out/r1z4_bench.json
