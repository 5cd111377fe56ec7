#!/bin/bash
# GPU check of the first-stage decoder: parity tests, then timing + per-op table
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_vae.py -m gpu -x -q > gpurun_out/vae_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/vae_tests.log
tail -25 gpurun_out/vae_tests.log
timeout 600 python tools/vae_profile.py ${1:-2} gpurun_out/vae_ops.csv > gpurun_out/vae_profile.log 2>&1; echo "rc=$?" >> gpurun_out/vae_profile.log
tail -40 gpurun_out/vae_profile.log
if [ -n "$2" ]; then
  timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/all_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/all_tests.log
  tail -5 gpurun_out/all_tests.log
fi
