#!/bin/bash
# Round-1 evidence on the final kernels: launch list of bench.py and full captures of the hot kernels
set -u
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_final.csv \
    python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_launches_final.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:probe_bitmap_kernel|build_kernel|compact_bitmap' -s 8 -c 10 \
    -f -o gpurun_out/prof_r1_final python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_full_final.log 2>&1
echo "ncu full rc=$?"
