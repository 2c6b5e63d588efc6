#!/bin/bash
set -u
mkdir -p gpurun_out
python tools/pcie_check.py > gpurun_out/r1u_pcie.txt 2>&1; cat gpurun_out/r1u_pcie.txt
RPT_TRACE=1 timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu --e2e-steps 3 > gpurun_out/r1u_bench.json 2> gpurun_out/r1u_bench.err
grep "rpt trace" gpurun_out/r1u_bench.err | tail -12
python -c "
import json; d=json.loads(open('gpurun_out/r1u_bench.json').read()); print(d['e2e'])"
nvidia-smi topo -m 2>/dev/null | head -8; lscpu | grep -E "NUMA|Model name|^CPU\(s\)" | head -8
