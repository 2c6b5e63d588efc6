#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "varchar or string or config3 or dictionary" > gpurun_out/r1r_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r1r_tests.log
tail -8 gpurun_out/r1r_tests.log
SWEEP_TYPE=varchar timeout 400 python tools/probe_sweep.py 2e8 3 8e4,1e7,5e7,1e8 > gpurun_out/r1r_varchar.txt 2>&1
rc=$?; cat gpurun_out/r1r_varchar.txt
[ $rc -eq 0 ] && SWEEP_TYPE=varchar timeout 600 ncu -k "regex:hash_varchar|probe_|build_kernel|compact_|scan_|count_seg" --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 40 --csv \
  --log-file gpurun_out/r1r_launches.csv python tools/probe_sweep.py 2e8 1 1e7 > gpurun_out/r1r_ncu.log 2>&1
echo "ncu rc=$?"
grep gpu__time gpurun_out/r1r_launches.csv | awk -F'","' '{print substr($5,1,60), $NF}' | tail -9
