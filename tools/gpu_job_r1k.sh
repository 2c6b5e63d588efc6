#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r1k_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r1k_tests.log
tail -5 gpurun_out/r1k_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests -m gpu -x -q -k "check_statistics or config1 or minmax_next" > gpurun_out/r1k_sanitizer.log 2>&1
echo "sanitizer rc=$?"; tail -4 gpurun_out/r1k_sanitizer.log
