#!/bin/bash
set -u
mkdir -p gpurun_out
export RPT_PROBE_PARTITION=1 RPT_PART_MINB=2
{
for st in 0 4 5 6 7; do RPT_PART_STORE=$st timeout 200 python tools/probe_sweep.py 1e9 5 1e8; done
} > gpurun_out/r1g4_sweep.txt 2>&1
cat gpurun_out/r1g4_sweep.txt
