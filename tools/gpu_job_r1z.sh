#!/bin/bash
# README-variant filter with shared-memory staging and block-range passes: tests, sweep, launch list, full capture
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_acero.py -m gpu -x -q > gpurun_out/r1z_acero_tests.log 2>&1
echo "acero tests rc=$?" >> gpurun_out/r1z_acero_tests.log
tail -15 gpurun_out/r1z_acero_tests.log
timeout 600 python tools/acero_sweep.py 1e9 5 8e4,1e7,1e8 1,16 > gpurun_out/r1z_acero_sweep.txt 2>&1
cat gpurun_out/r1z_acero_sweep.txt
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 300 --csv \
    --log-file gpurun_out/launches_r1z_acero.csv python tools/acero_sweep.py 1e9 1 1e8 16 > gpurun_out/r1z_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:abf_' -c 12 -f -o gpurun_out/prof_r1z_acero \
    python tools/acero_sweep.py 1e9 1 1e8 16 > gpurun_out/r1z_ncu_full.log 2>&1
echo "ncu full rc=$?"
