#!/bin/bash
set -u
mkdir -p gpurun_out
SWEEP_TYPE=varchar timeout 600 ncu -k "regex:hash_varchar|hash_entries|probe_|build_kernel|compact_|scan_|count_seg" --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 60 --csv \
  --log-file gpurun_out/r1q_launches.csv python tools/probe_sweep.py 2e8 1 1e7 > gpurun_out/r1q_ncu.log 2>&1
echo "ncu rc=$?"
grep gpu__time gpurun_out/r1q_launches.csv | awk -F'","' '{print substr($5,1,60), $NF}' | tail -24
