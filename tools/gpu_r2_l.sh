#!/bin/bash
# round-2 evidence (late): full GPU suite, full bench line, reference arm, smoke
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/r2_gpu_tests_v10.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_gpu_tests_v10.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_v8.json 2> gpurun_out/r2_bench_v8.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_v6.json 2> gpurun_out/r2_bench_ref_v6.err; echo "ref rc=$?"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python tools/route_check.py > gpurun_out/route_check_v3.txt 2>&1; grep -c "not the best" gpurun_out/route_check_v3.txt
