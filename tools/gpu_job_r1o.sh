#!/bin/bash
# config-3 shape: VARCHAR keys (decimal text), F = 128 MiB (n = 5e7) and 256 MiB, passes vs partition
set -u
mkdir -p gpurun_out
{
SWEEP_TYPE=varchar timeout 400 python tools/probe_sweep.py 2e8 3 1e7,5e7,1e8
SWEEP_TYPE=varchar RPT_PROBE_PARTITION=1 timeout 400 python tools/probe_sweep.py 2e8 3 5e7,1e8
SWEEP_TYPE=varchar RPT_PROBE_PARTITION=0 timeout 400 python tools/probe_sweep.py 2e8 3 5e7,1e8
} > gpurun_out/r1o_varchar.txt 2>&1
cat gpurun_out/r1o_varchar.txt
