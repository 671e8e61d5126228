#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/r2_gpu_tests_v4.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_gpu_tests_v4.log
tail -3 gpurun_out/r2_gpu_tests_v4.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_v2.json 2> gpurun_out/r2_bench_v2.err; echo "bench rc=$?"
tail -3 gpurun_out/r2_bench_v2.err
