#!/bin/bash
# verification after adopting the partition path for 4 slices: full GPU suite, sweeps (both key widths, passes vs partition), bench, ncu
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r1h_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r1h_tests.log
tail -4 gpurun_out/r1h_tests.log
{
echo "== default"
timeout 300 python tools/probe_sweep.py 1e9 5 8e4,1e7,5e7,1e8,3e8
echo "== RPT_PROBE_PARTITION=0 (sector-range passes)"
RPT_PROBE_PARTITION=0 timeout 300 python tools/probe_sweep.py 1e9 5 1e8,3e8
echo "== int32 default"
SWEEP_TYPE=int32 timeout 300 python tools/probe_sweep.py 1e9 5 1e8,3e8
echo "== int32 RPT_PROBE_PARTITION=0"
SWEEP_TYPE=int32 RPT_PROBE_PARTITION=0 timeout 300 python tools/probe_sweep.py 1e9 5 1e8,3e8
echo "== 50 % selectivity default / passes"
SWEEP_PPM=500000 timeout 300 python tools/probe_sweep.py 1e9 5 1e8
SWEEP_PPM=500000 RPT_PROBE_PARTITION=0 timeout 300 python tools/probe_sweep.py 1e9 5 1e8
} > gpurun_out/r1h_sweep.txt 2>&1
cat gpurun_out/r1h_sweep.txt
timeout 600 python bench.py > gpurun_out/r1h_bench.json 2> gpurun_out/r1h_bench.err
echo "bench rc=$?"; cat gpurun_out/r1h_bench.json
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/r1h_launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r1h_ncu.log 2>&1
echo "ncu rc=$?"
