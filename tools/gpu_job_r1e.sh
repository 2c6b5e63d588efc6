#!/bin/bash
# partitioned probe for filters beyond one L2 slice: parity, A/B timing, launch list
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_e.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu_e.log
{
python tools/probe_sweep.py 1e9 5 5e7,1e8,3e8
RPT_PROBE_PARTITION=0 python tools/probe_sweep.py 1e9 5 5e7,1e8,3e8
SWEEP_PPM=500000 python tools/probe_sweep.py 1e9 5 1e8
} > gpurun_out/sweep_e.txt 2>&1
cat gpurun_out/sweep_e.txt
python tools/probe_sweep.py 1e9 1 1e8 > gpurun_out/plain_sweep_e.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_e.csv \
    python tools/probe_sweep.py 1e9 1 1e8 > gpurun_out/ncu_launches_e.log 2>&1
echo "ncu rc=$?"
