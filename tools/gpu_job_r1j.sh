#!/bin/bash
# 8-GPU box: bench at N=8 and N=4 with the cheaper merge
set -u
mkdir -p gpurun_out
for n in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/r1j_bench_n$n.json 2> gpurun_out/r1j_bench_n$n.err
  echo "N=$n rc=$?"; python -c "import sys,json; d=json.loads(open('gpurun_out/r1j_bench_n$n.json').read()); print(d['value'], d['ms_per_step'], d['detail']['merge_ms'], d['detail']['build_ms'], d['detail']['probe_ms'], d['e2e']['value'], d['clocks'])"
done
