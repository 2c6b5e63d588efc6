#!/bin/bash
set -u
mkdir -p gpurun_out
SWEEP_TYPE=varchar timeout 600 ncu --set full --clock-control none --import-source on -k regex:hash_varchar_kernel -s 3 -c 1 \
  -o gpurun_out/varchar_r1s -f python tools/probe_sweep.py 2e8 1 1e7 > gpurun_out/r1s_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r1s_ncu.log
