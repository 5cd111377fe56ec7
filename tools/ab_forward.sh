#!/bin/bash
# A/B of the b=2 forward under environment variants: bash tools/ab_forward.sh "MOCA_X=0" "MOCA_X=1" ...
mkdir -p gpurun_out
i=0
for v in "$@"; do
  echo "== $v"
  env $v timeout 300 python tools/profile_forward.py 2 gpurun_out/ops_ab_$i.csv 2>&1 | grep -E "forward|sum of|gemm |attn|gn_|layernorm|convert"
  i=$((i+1))
done
