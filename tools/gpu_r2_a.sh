#!/bin/bash
# round-2 GPU call A: full GPU suite (incl. BASELINE-size parity), bench line with parity, ncu re-captures (cfg5 stream kernel, interp)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/r2_gpu_tests_v1.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_gpu_tests_v1.log
tail -3 gpurun_out/r2_gpu_tests_v1.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_v1.json 2> gpurun_out/r2_bench_v1.err; echo "bench rc=$?"
B="python bench.py --workload fftvcc --steps 1 --warmup 3 --no-cpu --no-e2e --no-others"
$B > gpurun_out/plain_fftvcc.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:fft_vcc -s 3 -c 1 -o gpurun_out/r2_fftvcc_stream $B > gpurun_out/ncu_fftvcc.log 2>&1
python tools/interp_check.py > gpurun_out/r2_interp_check.txt 2>&1 && ncu --set full --clock-control none --import-source on -k regex:interp_fir -s 3 -c 1 -o gpurun_out/r2_interp python tools/interp_check.py > gpurun_out/ncu_interp.log 2>&1
cat gpurun_out/r2_interp_check.txt
ls -la gpurun_out | tail -12
