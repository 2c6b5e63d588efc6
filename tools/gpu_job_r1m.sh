#!/bin/bash
set -u
mkdir -p gpurun_out
{
for mib in 32 64 86 128 0; do RPT_SLICE_MIB=$mib timeout 200 python tools/probe_sweep.py 2e8 5 1e8,3e8; done
} > gpurun_out/r1m_build_slices.txt 2>&1
cat gpurun_out/r1m_build_slices.txt
