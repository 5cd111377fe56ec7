#!/bin/bash
# Runs each section of tools/gpu_check.py in its own process (a sticky CUDA fault or a hang in one section
# must not take the others down); logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.log 2>&1
for s in "$@"; do
  timeout 420 python tools/gpu_check.py $s > gpurun_out/check_$s.log 2>&1
  echo "--- $s exit $?" | tee -a gpurun_out/check_$s.log
  tail -n 60 gpurun_out/check_$s.log
done
