#!/bin/bash
mkdir -p gpurun_out
export MOCA_NO_GRAPH=1
python tools/one_forward.py 1 || exit 1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k 'regex:gemm_tc_kernel<\(int\)160, \(int\)0' -c 14 -o gpurun_out/ncu_gemm160 -f python tools/one_forward.py 1 > gpurun_out/ncu_gemm160.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_gemm160.log; ls -la gpurun_out/*.ncu-rep
