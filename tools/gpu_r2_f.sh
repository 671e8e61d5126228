#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_interp_fir.py tests/test_gpu_rational_resampler.py -m gpu -q > gpurun_out/r2_interp_tests.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2_interp_tests.log
python tools/interp_check.py 2>&1 | tee gpurun_out/r2_interp_check_v2.txt
python tools/resamp_check.py 2>&1 | tee gpurun_out/r2_resamp_check_v1.txt
