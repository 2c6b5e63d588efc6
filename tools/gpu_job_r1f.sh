#!/bin/bash
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "beyond_one or config2" > gpurun_out/pytest_gpu_f.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_f.log
python tools/probe_sweep.py 1e9 5 5e7,1e8,3e8 2>&1 | tee gpurun_out/sweep_f.txt
python tools/probe_sweep.py 1e9 1 1e8 > gpurun_out/plain_sweep_f.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:probe_partition_kernel|probe_bucket_kernel' -c 2 \
    -f -o gpurun_out/prof_r1f python tools/probe_sweep.py 1e9 1 1e8 > gpurun_out/ncu_full_f.log 2>&1
echo "ncu rc=$?"
