#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r1t_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r1t_tests.log
tail -5 gpurun_out/r1t_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/r1t_bench.json 2> gpurun_out/r1t_bench.err
echo "bench rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/r1t_bench.json').read())
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['cpu_baseline']['value'], d['gpu_launches'], d['clocks'])"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1t_ref.json 2> gpurun_out/r1t_ref.err
echo "ref rc=$?"; cut -c1-300 gpurun_out/r1t_ref.json
