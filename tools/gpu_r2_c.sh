#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_staging.py -m gpu -q > gpurun_out/r2_staging_v6.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_staging_v6.log
( time python bench.py --steps 10 --warmup 3 --no-others --no-cpu > gpurun_out/r2_bench_e2e_v3.json 2> gpurun_out/r2_bench_e2e_v3.err ) 2>&1 | tail -3; echo "bench rc=$?"
tail -5 gpurun_out/r2_bench_e2e_v3.err
