#!/bin/bash
set -u
mkdir -p gpurun_out
export RPT_PROBE_PARTITION=1 RPT_PART_MINB=2
timeout 600 ncu --set full --clock-control none --import-source on -k regex:probe_partition_kernel -c 1 \
  -o gpurun_out/part_r1g -f python tools/probe_sweep.py 1e9 1 1e8 > gpurun_out/r1g_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r1g_ncu.log; ls -la gpurun_out/*.ncu-rep
