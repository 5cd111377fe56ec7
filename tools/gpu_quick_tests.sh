#!/bin/bash
# run a pytest selection on the GPU box: tools/gpu_quick_tests.sh <pytest args>
mkdir -p gpurun_out
timeout 1500 python -m pytest "$@" -m gpu -x -q > gpurun_out/quick_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/quick_tests.log
tail -30 gpurun_out/quick_tests.log
