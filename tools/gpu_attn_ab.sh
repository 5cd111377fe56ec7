#!/bin/bash
# attention: unit parity, micro-bench, and forward A/B of the working tree's library against the committed one (gpurun_tmp/base.so)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k "attention" 2>&1 | tail -2
timeout 300 python tools/attn_bench.py 2>&1 | head -5
for i in 1 2; do timeout 300 python tools/profile_forward.py 2 gpurun_out/ops_new.csv 2>&1 | grep -E "forward|attn_self|attn_cross"; done
