#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r1l_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r1l_tests.log
tail -25 gpurun_out/r1l_tests.log
