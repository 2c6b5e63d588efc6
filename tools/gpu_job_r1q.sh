#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "varchar or string or config3 or dictionary" > gpurun_out/r1q_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r1q_tests.log
tail -4 gpurun_out/r1q_tests.log
SWEEP_TYPE=varchar timeout 400 python tools/probe_sweep.py 2e8 3 1e7,5e7 > gpurun_out/r1q_varchar.txt 2>&1
rc=$?; cat gpurun_out/r1q_varchar.txt
[ $rc -eq 0 ] && SWEEP_TYPE=varchar timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 60 --csv \
  --log-file gpurun_out/r1q_launches.csv python tools/probe_sweep.py 2e8 1 1e7 > gpurun_out/r1q_ncu.log 2>&1
echo "ncu rc=$?"
grep -E "hash_varchar|probe_|build_kernel|compact" gpurun_out/r1q_launches.csv | grep gpu__time | awk -F'","' '{print $5, $NF}' | tail -14
