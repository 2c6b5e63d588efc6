#!/bin/bash
# README-variant filter, final round-1 kernels: tests, sweep, launch list, full captures per kernel family; then the
# whole GPU suite once more on the shipped library
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_acero.py -m gpu -q > gpurun_out/r1z2_acero_tests.log 2>&1
echo "acero tests rc=$?" >> gpurun_out/r1z2_acero_tests.log
tail -15 gpurun_out/r1z2_acero_tests.log
timeout 600 python tools/acero_sweep.py 1e9 5 2e4,8e4,1e7,1e8 1,16 > gpurun_out/r1z2_acero_sweep.txt 2>&1
cat gpurun_out/r1z2_acero_sweep.txt
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 300 --csv \
    --log-file gpurun_out/launches_r1z2_acero.csv python tools/acero_sweep.py 1e9 1 1e8 1,16 > gpurun_out/r1z2_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:abf_probe' -c 4 -f -o gpurun_out/prof_r1z2_acero_probe \
    python tools/acero_sweep.py 1e9 1 1e8 1 > gpurun_out/r1z2_ncu_full_probe.log 2>&1
echo "ncu full probe rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:abf_build|abf_fold|abf_popcount' -c 12 -f -o gpurun_out/prof_r1z2_acero_build \
    python tools/acero_sweep.py 1e8 1 1e8 4 > gpurun_out/r1z2_ncu_full_build.log 2>&1
echo "ncu full build rc=$?"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r1z2_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r1z2_tests.log
tail -4 gpurun_out/r1z2_tests.log
