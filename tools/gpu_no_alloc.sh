#!/bin/bash
# the allocation-free entry point's test alone in a fresh process, then the whole GPU suite (it runs last there)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/no_alloc_alone.log 2>&1; echo "pytest rc=$?" >> gpurun_out/no_alloc_alone.log
tail -15 gpurun_out/no_alloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/all_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/all_tests.log
tail -8 gpurun_out/all_tests.log
