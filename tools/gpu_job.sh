#!/bin/bash
# One parametrised GPU job script (replaces the per-run gpu_job_r1*.sh files of round 1):
#   gpurun --timeout N -- 'bash tools/gpu_job.sh <job> [args]'
# Every job writes into gpurun_out/ (merged back by gpurun) and guards risky steps with its own timeout.
set -u
mkdir -p gpurun_out
job=${1:-help}
shift || true
case "$job" in
part_lab)   # prototype: one-pass partition into shared-memory-sized buckets + shared-memory lookup
	out=gpurun_out/part_lab_${1:-r2a}.txt
	: > $out
	timeout -s KILL 120 tools/part_lab atoms >> $out 2>&1
	# sectors block kpt rows log2_chunk density fake
	for cfg in "25 1024 8 1000000000 32 1 1" "25 512 4 1000000000 32 1 1" "25 512 4 1000000000 32 1 0" "25 256 8 1000000000 32 1 0" \
	           "25 1024 8 1000000000 32 2 0" "25 1024 8 1000000000 27 1 0" "25 1024 8 1000000000 26 1 0" "25 1024 8 1000000000 27 0 0"; do
		timeout -s KILL 90 tools/part_lab part $cfg >> $out 2>&1
		echo "rc=$?" >> $out
	done
	cat $out
	;;
part2)      # v2 prototype sweep
	out=gpurun_out/part2_${1:-r2d}.txt
	: > $out
	# sectors block kpt rows log2_chunk density slots skew
	for cfg in "25 1024 4 1000000000 26 1 8 0" "25 1024 2 1000000000 26 1 8 0" "25 768 4 1000000000 26 1 8 0" "25 512 8 1000000000 26 1 8 0" \
	           "25 1024 4 1000000000 26 0 8 0" "25 1024 4 1000000000 26 2 8 0" "25 1024 4 1000000000 25 1 8 0" \
	           "24 1024 4 1000000000 26 1 16 0" "24 1024 4 1000000000 26 1 8 0" "25 1024 4 200000000 26 1 8 4" "25 1024 4 200000000 26 1 8 14"; do
		timeout -s KILL 90 tools/part_lab part2 $cfg >> $out 2>&1
		echo "rc=$?" >> $out
	done
	cat $out
	;;
part3)      # v3 prototype sweep
	out=gpurun_out/part3_${1:-r2e}.txt
	: > $out
	# sectors block kpp rows log2_chunk density slots skew
	for cfg in "25 1024 4 1000000000 27 1 8 0" "25 1024 2 1000000000 27 1 8 0" "25 1024 8 1000000000 27 1 8 0" "25 768 4 1000000000 27 1 8 0" \
	           "25 1024 4 1000000000 27 0 8 0" "25 1024 4 1000000000 26 1 8 0" \
	           "24 1024 4 1000000000 27 1 16 0" "24 1024 4 1000000000 27 1 8 0" "25 1024 4 200000000 27 1 8 4" "25 1024 4 200000000 27 1 8 14"; do
		timeout -s KILL 90 tools/part_lab part3 $cfg >> $out 2>&1
		echo "rc=$?" >> $out
	done
	cat $out
	;;
part4)      # v4 prototype sweep
	out=gpurun_out/part4_${1:-r2f}.txt
	: > $out
	# sectors block kpp rows log2_chunk density slots skew
	for cfg in "25 1024 4 1000000000 27 1 8 0" "25 1024 2 1000000000 27 1 8 0" "25 1024 8 1000000000 27 1 8 0" "25 768 4 1000000000 27 1 8 0" \
	           "25 1024 4 1000000000 27 0 8 0" "25 1024 4 1000000000 26 1 8 0" \
	           "24 1024 4 1000000000 27 1 16 0" "24 1024 4 1000000000 27 1 8 0" "25 1024 4 200000000 27 1 8 4" "25 1024 4 200000000 27 1 8 14" \
	           "25 1024 4 16777216 27 1 8 0"; do
		timeout -s KILL 90 tools/part_lab part4 $cfg >> $out 2>&1
		echo "rc=$?" >> $out
	done
	cat $out
	;;
ncu_py)     # full ncu capture of product kernels: ncu_py <tag> <kernel regex> <skip> <count> <python script + args...>
	tag=$1; rx=$2; skip=$3; cnt=$4; shift 4
	timeout 300 python "$@" > gpurun_out/ncu_${tag}_plain.log 2>&1 &&
	timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt \
		-o gpurun_out/ncu_${tag} -f python "$@" > gpurun_out/ncu_${tag}.log 2>&1
	echo "rc=$?"; tail -3 gpurun_out/ncu_${tag}.log
	;;
ncu_lab)    # full ncu capture of one prototype kernel: ncu_lab <tag> <kernel regex> <part_lab args...>
	tag=$1; rx=$2; shift 2
	timeout -s KILL 120 tools/part_lab "$@" > gpurun_out/ncu_lab_${tag}_plain.log 2>&1 &&
	timeout -s KILL 900 ncu --set full --clock-control none --import-source on -k regex:$rx -c 1 \
		-o gpurun_out/ncu_lab_${tag} -f tools/part_lab "$@" > gpurun_out/ncu_lab_${tag}.log 2>&1
	echo "rc=$?"; tail -3 gpurun_out/ncu_lab_${tag}.log
	;;
e2e)        # configs 1 / 4 / 5 through DuckDB + the GPU-backed reference extension, diffed against plain DuckDB
	tag=${1:-r2}
	timeout 900 python integration/e2e/run_e2e.py --ext gpu > gpurun_out/e2e_${tag}_default.jsonl 2>&1
	echo "e2e default rc=$?"; cat gpurun_out/e2e_${tag}_default.jsonl | cut -c1-600
	timeout 1500 python integration/e2e/run_e2e.py --ext gpu --tpch-sf ${2:-10} --job-scale ${3:-4} --live-baseline --only config > gpurun_out/e2e_${tag}_large.jsonl 2>&1
	echo "e2e large rc=$?"; cat gpurun_out/e2e_${tag}_large.jsonl | cut -c1-600
	timeout 900 python integration/e2e/run_e2e.py --ext stock --tpch-sf ${2:-10} --job-scale ${3:-4} --out gpurun_out/stock > gpurun_out/e2e_${tag}_large_stock.jsonl 2>&1
	cat gpurun_out/e2e_${tag}_large_stock.jsonl | cut -c1-300
	;;
pcie)       # raw pinned H2D/D2H with N concurrent processes: pcie <tag> <maxN>
	tag=$1; maxn=$2
	: > gpurun_out/pcie_matrix_${tag}.jsonl
	for n in 1 2 4 8; do
		[ $n -le $maxn ] || continue
		timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29521 \
			tools/pcie_matrix.py 2>/dev/null | grep "^{" >> gpurun_out/pcie_matrix_${tag}.jsonl
	done
	nvidia-smi topo -m > gpurun_out/topo_${tag}.txt 2>&1; lscpu | head -25 >> gpurun_out/topo_${tag}.txt
	cat gpurun_out/pcie_matrix_${tag}.jsonl | cut -c1-900
	;;
benchN)     # multi-GPU bench under torchrun: benchN <tag> <N> [extra bench args]
	tag=$1; n=$2; shift 2
	timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
		bench.py --gpus $n --no-cpu "$@" > gpurun_out/bench_${tag}_n${n}.json 2> gpurun_out/bench_${tag}_n${n}.err
	echo "benchN rc=$?"; cut -c1-300 gpurun_out/bench_${tag}_n${n}.json; tail -5 gpurun_out/bench_${tag}_n${n}.err
	python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_${tag}_n${n}.json").read())
    print({k: d[k] for k in ("value", "ms_per_step", "n_gpus")}, d["detail"]["build_ms"], d["detail"]["merge_ms"], d["detail"]["probe_ms"], d["detail"]["parity"], d.get("e2e", {}).get("value"), d["detail"].get("sustained", {}).get("value"))
except Exception as e:
    print("no line:", e)
PY
	;;
partvar)    # partitioned-probe variants (32 MiB slices, slice 0 as a bucket): partvar <tag>
	out=gpurun_out/partvar_${1:-r2x}.txt
	: > $out
	for knobs in "" "RPT_PART_SLICE_MIB=32" "RPT_PART_INLINE0=0" "RPT_PART_SLICE_MIB=32 RPT_PART_INLINE0=0"; do
		echo "== knobs: ${knobs:-default}" >> $out
		env $knobs timeout -s KILL 300 python tools/probe_sweep.py 1e9 5 1e8 >> $out 2>&1
		echo "rc=$?" >> $out
		env $knobs SWEEP_PPM=500000 timeout -s KILL 300 python tools/probe_sweep.py 1e9 5 1e8 >> $out 2>&1
		echo "rc=$?" >> $out
	done
	cat $out
	;;
partpipe)   # chunk-pipelined partitioned probe (code as of commit 0fbcbed, removed afterwards): partpipe <tag>
	out=gpurun_out/partpipe_${1:-r2y}.txt
	: > $out
	for knobs in "" "RPT_PART_PIPE=8" "RPT_PART_PIPE=8 RPT_PART_INLINE0=0" "RPT_PART_PIPE=8 RPT_PART_INLINE0=0 RPT_PART_PIPE_BCTAS=2" \
	             "RPT_PART_PIPE=8 RPT_PART_INLINE0=0 RPT_PART_PIPE_BCTAS=4" "RPT_PART_PIPE=4 RPT_PART_INLINE0=0" "RPT_PART_PIPE=16 RPT_PART_INLINE0=0" \
	             "RPT_PART_PIPE=8 RPT_PART_PIPE_BCTAS=2" "RPT_PART_PIPE=16" "RPT_PART_PIPE=8 RPT_PART_PIPE_PCTAS=2 RPT_PART_PIPE_BCTAS=6"; do
		echo "== knobs: ${knobs:-default}" >> $out
		env $knobs timeout -s KILL 300 python tools/probe_sweep.py 1e9 5 1e8 >> $out 2>&1
		echo "rc=$?" >> $out
	done
	cat $out
	;;
prefilter)  # folded pre-filter: the whole GPU suite, then on/off sweeps at config 2 (1 %, 10 %, 50 %), 512 MiB, other key types: prefilter <tag>
	tag=${1:-r2p}
	timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/tests_${tag}.log 2>&1
	echo "tests rc=$?" | tee -a gpurun_out/tests_${tag}.log
	tail -5 gpurun_out/tests_${tag}.log
	out=gpurun_out/prefilter_${tag}.txt
	: > $out
	for knobs in "" "RPT_PREFILTER=0"; do
		for ppm in 100000 10000 500000; do
			echo "== knobs: ${knobs:-default} ppm=$ppm" >> $out
			env $knobs SWEEP_PPM=$ppm timeout -s KILL 300 python tools/probe_sweep.py 1e9 5 1e8 >> $out 2>&1
			echo "rc=$?" >> $out
		done
	done
	for knobs in "" "RPT_PREFILTER=0"; do
		echo "== knobs: ${knobs:-default} 512 MiB" >> $out
		env $knobs timeout -s KILL 300 python tools/probe_sweep.py 1e9 3 1.8e8,3e8 >> $out 2>&1
		echo "rc=$?" >> $out
	done
	for typ in two int32 varchar; do
		for knobs in "" "RPT_PREFILTER=0"; do
			echo "== knobs: ${knobs:-default} type=$typ" >> $out
			env $knobs SWEEP_TYPE=$typ timeout -s KILL 400 python tools/probe_sweep.py 4e8 3 1e8 >> $out 2>&1
			echo "rc=$?" >> $out
		done
	done
	grep -v "^rc=0\|^# m=" $out
	cmd="python tools/probe_sweep.py 1e9 1 1e8"
	timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 \
		--csv --log-file gpurun_out/launches_prefilter_${tag}.csv $cmd > gpurun_out/ncu_prefilter_${tag}.log 2>&1
	echo "ncu rc=$?"
	python tools/launch_summary.py gpurun_out/launches_prefilter_${tag}.csv > gpurun_out/launches_prefilter_${tag}.md 2>&1
	head -30 gpurun_out/launches_prefilter_${tag}.md
	;;
tests)      # the GPU parity suite
	timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/tests_${1:-r2}.log 2>&1
	echo "tests rc=$?" >> gpurun_out/tests_${1:-r2}.log
	tail -15 gpurun_out/tests_${1:-r2}.log
	;;
bench)      # default bench line (+ reference arm)
	tag=${1:-r2}
	timeout 900 python bench.py > gpurun_out/bench_${tag}.json 2> gpurun_out/bench_${tag}.err
	echo "bench rc=$?"
	cut -c1-600 gpurun_out/bench_${tag}.json
	tail -5 gpurun_out/bench_${tag}.err
	;;
evidence)   # round-end evidence: tests, both bench arms, then the ncu launch list of the same bench command
	tag=${1:-r2}
	timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/tests_${tag}.log 2>&1
	echo "tests rc=$?" | tee -a gpurun_out/tests_${tag}.log
	tail -3 gpurun_out/tests_${tag}.log
	timeout 900 python bench.py --impl reference > gpurun_out/bench_${tag}_reference.json 2> gpurun_out/bench_${tag}_reference.err
	echo "reference arm rc=$?"
	timeout 900 python bench.py > gpurun_out/bench_${tag}.json 2> gpurun_out/bench_${tag}.err
	echo "bench rc=$?"
	cut -c1-400 gpurun_out/bench_${tag}.json
	cmd="python bench.py --steps 2 --warmup 3 --no-sweep --no-config3 --no-cpu --no-e2e --sustain-s 0"
	timeout 600 $cmd > gpurun_out/bench_${tag}_short.json 2> gpurun_out/bench_${tag}_short.err
	rc=$?
	echo "short bench rc=$rc"
	if [ $rc -eq 0 ]; then
		timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 \
			--csv --log-file gpurun_out/launches_${tag}.csv $cmd > gpurun_out/ncu_${tag}.log 2>&1
		echo "ncu rc=$?"
		python tools/launch_summary.py gpurun_out/launches_${tag}.csv > gpurun_out/launches_${tag}.md 2>&1
		head -30 gpurun_out/launches_${tag}.md
	fi
	;;
*)
	echo "jobs: part_lab tests bench evidence"
	;;
esac
