#!/bin/bash
# Quick GPU visit: UNet/ops parity tests + per-op profile of the b=2 forward.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 12 gpurun_out/pytest_gpu.log
timeout 600 python tools/profile_forward.py 2 gpurun_out/profile_ops_b2.csv 2>&1 | tail -16
