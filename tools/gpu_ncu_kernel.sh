#!/bin/bash
# ncu --set full of selected kernels of one eager b=2 forward: bash tools/gpu_ncu_kernel.sh <name> '<regex>' <skip> <count>
mkdir -p gpurun_out
export MOCA_NO_GRAPH=1
python tools/one_forward.py 1 && \
ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k "regex:$2" -s $3 -c $4 -o gpurun_out/ncu_$1 -f python tools/one_forward.py 1 > gpurun_out/ncu_$1.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_$1.log; ls -la gpurun_out/ncu_$1.ncu-rep
