#!/bin/bash
set -u
mkdir -p gpurun_out
export RPT_PROBE_PARTITION=1 RPT_PART_MINB=2
{
for st in 0 1 2 3; do RPT_PART_STORE=$st timeout 200 python tools/probe_sweep.py 1e9 5 1e8; done
for mib in 48 64 80 96; do RPT_L2_PERSIST_MIB=$mib timeout 200 python tools/probe_sweep.py 1e9 5 1e8; done
RPT_L2_PERSIST_MIB=80 RPT_PART_STORE=1 timeout 200 python tools/probe_sweep.py 1e9 5 1e8
} > gpurun_out/r1g3_sweep.txt 2>&1
cat gpurun_out/r1g3_sweep.txt
