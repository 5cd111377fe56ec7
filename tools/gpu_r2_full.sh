#!/bin/bash
# Round-2 evidence visit (1 GPU): full bench line, launch list + DRAM traffic of the timed region, per-op profile,
# ncu --set full of the dominant kernels.
mkdir -p gpurun_out
timeout 1200 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/bench.json; tail -n 3 gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_reference.json 2>> gpurun_out/bench.err; echo "reference rc=$?"; cut -c1-300 gpurun_out/bench_reference.json
timeout 600 python tools/profile_forward.py 2 gpurun_out/profile_ops_b2.csv 2>&1 | tail -12
# launch list of the timed region only (bench.py brackets it with cudaProfilerStart/Stop when MOCA_CUDA_PROFILER=1)
python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/bench_short.json 2>&1 && \
MOCA_CUDA_PROFILER=1 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2200 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/ncu_bench.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/launches.csv
python tools/summarize_launches.py gpurun_out/launches.csv 2 > gpurun_out/launches_summary.txt 2>&1; tail -n 3 gpurun_out/launches_summary.txt | cut -c1-300
export MOCA_NO_GRAPH=1
python tools/one_forward.py 1 > /dev/null 2>&1 && {
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:gemm_tc_kernel<\(int\)160, \(int\)0, \(bool\)1, \(bool\)0' -s 8 -c 2 -o gpurun_out/ncu_conv160 -f python tools/one_forward.py 1 > gpurun_out/ncu_conv160.log 2>&1; echo "ncu conv160 rc=$?"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:attn_tc2_kernel' -s 1 -c 2 -o gpurun_out/ncu_attn2 -f python tools/one_forward.py 1 > gpurun_out/ncu_attn2.log 2>&1; echo "ncu attn2 rc=$?"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:gn_apply_ab_kernel' -s 6 -c 4 -o gpurun_out/ncu_gnapply -f python tools/one_forward.py 1 > gpurun_out/ncu_gnapply.log 2>&1; echo "ncu gn_apply rc=$?"
}
ls -la gpurun_out/*.ncu-rep
