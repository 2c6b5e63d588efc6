#!/bin/bash
# Round-1 evidence on the final kernels: full capture of one bench step's hot kernels
set -u
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/r1n_plain_bench.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on \
    -k 'regex:probe_partition_kernel|probe_bucket_kernel|build_kernel|compact_bitmap_kernel|count_segments_kernel' -s 10 -c 10 \
    -f -o gpurun_out/prof_r1n python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/r1n_ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/r1n_ncu_full.log; ls -la gpurun_out/prof_r1n.ncu-rep
