#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:probe_bitmap_kernel -s 2 -c 1 \
  -o gpurun_out/smem_r1w -f python tools/probe_sweep.py 1e9 1 8e4 > gpurun_out/r1w_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r1w_ncu.log
