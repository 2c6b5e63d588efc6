#!/bin/bash
# README-variant filter with cache-policy hints: tests, sweep, full capture of the probe; N=2 bench of the main path
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_acero.py -m gpu -q > gpurun_out/r1z3_acero_tests.log 2>&1
echo "acero tests rc=$?" >> gpurun_out/r1z3_acero_tests.log
tail -5 gpurun_out/r1z3_acero_tests.log
timeout 600 python tools/acero_sweep.py 1e9 5 8e4,1e7,1e8 1,16 > gpurun_out/r1z3_acero_sweep.txt 2>&1
cat gpurun_out/r1z3_acero_sweep.txt
ncu --set full --clock-control none --import-source on -k 'regex:abf_probe|abf_build' -c 6 -f -o gpurun_out/prof_r1z3_acero \
    python tools/acero_sweep.py 1e9 1 1e8 1 > gpurun_out/r1z3_ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r1z3_bench_n2.json 2> gpurun_out/r1z3_bench_n2.err
echo "bench n2 rc=$?"
cut -c1-600 gpurun_out/r1z3_bench_n2.json
