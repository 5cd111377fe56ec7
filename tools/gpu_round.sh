#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench, steady-state launch list, per-op profile.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -n 5 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -n 5 gpurun_out/bench.err
# launch list of the timed region only (bench.py brackets it with cudaProfilerStart/Stop when MOCA_CUDA_PROFILER=1)
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.json 2>&1 && \
MOCA_CUDA_PROFILER=1 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2200 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
echo "ncu rc=$?"; tail -n 1 gpurun_out/ncu_bench.log | cut -c1-200; wc -l gpurun_out/launches.csv
timeout 600 python tools/profile_forward.py 2 gpurun_out/profile_ops_b2.csv 2>&1 | tail -12
