#!/bin/bash
# compaction v2 + 3-stream host pipeline: parity, sweeps, e2e trace, ncu launch list + full capture
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_c.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_c.log
{
python tools/probe_sweep.py 1e9 5
SWEEP_PPM=10000 python tools/probe_sweep.py 1e9 5 1e7,1e8
SWEEP_PPM=500000 python tools/probe_sweep.py 1e9 5 1e7,1e8
} > gpurun_out/sweep_c.txt 2>&1
cat gpurun_out/sweep_c.txt
for rows in 16777216 4194304 67108864; do
  echo "== RPT_CHUNK_ROWS=$rows"
  RPT_CHUNK_ROWS=$rows RPT_TRACE=1 python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 2 2> gpurun_out/e2e_trace_$rows.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('value', d['value'], 'e2e', d['e2e'])"
  cat gpurun_out/e2e_trace_$rows.err | tail -8
done 2>&1 | tee gpurun_out/e2e_c.txt
python tools/probe_sweep.py 1e9 1 1e7,1e8 > gpurun_out/plain_sweep.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches_c.csv \
    python tools/probe_sweep.py 1e9 1 1e7,1e8 > gpurun_out/ncu_launches_c.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:probe_bitmap_kernel|compact_bitmap' -s 10 -c 6 \
    -f -o gpurun_out/prof_r1c python tools/probe_sweep.py 1e9 1 1e7,1e8 > gpurun_out/ncu_full_c.log 2>&1
echo "ncu full rc=$?"
