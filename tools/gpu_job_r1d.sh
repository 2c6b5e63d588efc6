#!/bin/bash
# compaction v3 (shared-memory runs): parity, sweeps at 1/10/50 %, launch list
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_d.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_d.log
{
python tools/probe_sweep.py 1e9 5
SWEEP_PPM=10000 python tools/probe_sweep.py 1e9 5 1e7,1e8
SWEEP_PPM=500000 python tools/probe_sweep.py 1e9 5 1e7,1e8
} > gpurun_out/sweep_d.txt 2>&1
cat gpurun_out/sweep_d.txt
python tools/probe_sweep.py 1e9 1 1e7 > gpurun_out/plain_sweep_d.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 30 --csv --log-file gpurun_out/launches_d.csv \
    python tools/probe_sweep.py 1e9 1 1e7 > gpurun_out/ncu_launches_d.log 2>&1
grep -E "compact|probe_bitmap" gpurun_out/launches_d.csv | awk -F'","' '{print $5, $NF}' | cut -c1-160 | head -8
