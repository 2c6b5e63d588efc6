#!/bin/bash
# Round-1 GPU job: parity suite, bench (with sweep), memory-system microbenchmark, ncu launch list and
# one full capture of the hot kernels.  Run as: gpurun --timeout 1800 -- 'bash tools/gpu_job_r1.sh'
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks.mem,power.draw,memory.total --format=csv > gpurun_out/gpu_info.csv
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
python bench.py --sweep > gpurun_out/bench_sweep.json 2> gpurun_out/bench_sweep.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench_sweep.json
timeout 600 tools/mem_lab 30 > gpurun_out/mem_lab.txt 2>&1; echo "mem_lab rc=$?"
python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/plain_launches.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
python tools/probe_case.py 1e8 1e9 100000 1 > gpurun_out/plain_probe_case.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:probe_pass_kernel|build_kernel|compact_kernel' -c 9 \
    -f -o gpurun_out/prof_r1 python tools/probe_case.py 1e8 1e9 100000 1 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
cat gpurun_out/plain_probe_case.log
