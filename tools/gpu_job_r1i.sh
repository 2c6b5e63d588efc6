#!/bin/bash
# 2-GPU: bench with the cheaper merge (flag in the first barrier, last barrier deferred), both merge modes
set -u
mkdir -p gpurun_out
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 "$@"; }
run --steps 10 --warmup 3 > gpurun_out/r1i_bench_n2.json 2> gpurun_out/r1i_bench_n2.err
echo "rc=$?"; cat gpurun_out/r1i_bench_n2.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['detail']['merge_ms'], d['detail']['build_ms'], d['detail']['probe_ms'], d['e2e']['value'])"
run --steps 10 --warmup 3 --merge allgather > gpurun_out/r1i_bench_n2_ag.json 2> gpurun_out/r1i_bench_n2_ag.err
echo "rc=$?"; cat gpurun_out/r1i_bench_n2_ag.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['detail']['merge_ms'])"
tail -3 gpurun_out/r1i_bench_n2.err
