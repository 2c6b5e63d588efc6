#!/bin/bash
# lean partition kernel: parity with the partition forced on, then sweeps (default passes vs partition, 3 vs 2 CTAs/SM)
set -u
mkdir -p gpurun_out
RPT_PROBE_PARTITION=1 timeout 900 python -m pytest tests -m gpu -x -q -k "flat_columns or beyond_one or config2 or skew" > gpurun_out/r1g_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r1g_tests.log
tail -5 gpurun_out/r1g_tests.log
{
echo "== default (sector-range passes at 2/4 slices)"
timeout 300 python tools/probe_sweep.py 1e9 5 5e7,1e8,3e8
echo "== RPT_PROBE_PARTITION=1 (3 CTAs/SM where they fit)"
RPT_PROBE_PARTITION=1 timeout 300 python tools/probe_sweep.py 1e9 5 5e7,1e8,3e8
echo "== RPT_PROBE_PARTITION=1 RPT_PART_MINB=2"
RPT_PROBE_PARTITION=1 RPT_PART_MINB=2 timeout 300 python tools/probe_sweep.py 1e9 5 5e7,1e8,3e8
} > gpurun_out/r1g_sweep.txt 2>&1
cat gpurun_out/r1g_sweep.txt
