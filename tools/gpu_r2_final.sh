#!/bin/bash
# Final round-2 evidence visit (1 GPU): GPU test suite, smoke, full bench line + reference arm, per-op profile, launch list with DRAM
# traffic of the timed region.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/all_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/all_tests.log; tail -3 gpurun_out/all_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/smoke.log
timeout 1200 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-900 gpurun_out/bench.json; tail -n 3 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_reference.json 2>> gpurun_out/bench.err; echo "reference rc=$?"; cut -c1-300 gpurun_out/bench_reference.json
timeout 600 python tools/profile_forward.py 2 gpurun_out/profile_ops_b2.csv 2>&1 | tail -12
python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/bench_short.json 2>&1 && \
MOCA_CUDA_PROFILER=1 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2200 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/ncu_bench.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/launches.csv
python tools/summarize_launches.py gpurun_out/launches.csv 2 > gpurun_out/launches_summary.txt 2>&1; tail -n 3 gpurun_out/launches_summary.txt | cut -c1-300
