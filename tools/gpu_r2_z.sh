#!/bin/bash
# late round-2 evidence: full GPU suite, bench, reference arm, smoke, launch list, ncu of the two new kernels
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/r2_gpu_tests_v15.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_gpu_tests_v15.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_v13.json 2> gpurun_out/r2_bench_v13.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_v11.json 2> gpurun_out/r2_bench_ref_v11.err; echo "ref rc=$?"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e"
$B > gpurun_out/plain_all.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2_launches_bench_v7.csv $B > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
timeout 900 python tools/route_check.py D=1 D=2 D=4 D=8 D=10 D=16 D=20 D=25 D=32 > gpurun_out/route_check_v9.txt 2>&1; grep -c "not the best" gpurun_out/route_check_v9.txt; true
