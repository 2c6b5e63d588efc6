#!/bin/bash
# Re-entry check of the restored tree + the README-variant filter: full GPU suite, smoke, both bench arms, sweep
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_acero.py -m gpu -x -q > gpurun_out/r1y_acero_tests.log 2>&1
echo "acero tests rc=$?" >> gpurun_out/r1y_acero_tests.log
tail -15 gpurun_out/r1y_acero_tests.log
timeout 600 python tools/acero_sweep.py 1e9 5 8e4,1e7,1e8 1,16 > gpurun_out/r1y_acero_sweep.txt 2>&1
cat gpurun_out/r1y_acero_sweep.txt
timeout 1500 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_acero.py > gpurun_out/r1y_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r1y_tests.log
tail -4 gpurun_out/r1y_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r1y_smoke.log 2>&1
tail -2 gpurun_out/r1y_smoke.log
timeout 600 python bench.py > gpurun_out/r1y_bench.json 2> gpurun_out/r1y_bench.err
cat gpurun_out/r1y_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1y_bench_ref.json 2> gpurun_out/r1y_bench_ref.err
cat gpurun_out/r1y_bench_ref.json
