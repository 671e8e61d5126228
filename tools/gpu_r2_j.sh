#!/bin/bash
# round-2 evidence (late): full GPU suite, full bench line, reference arm, smoke
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/r2_gpu_tests_v9.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_gpu_tests_v9.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_v7.json 2> gpurun_out/r2_bench_v7.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_v5.json 2> gpurun_out/r2_bench_ref_v5.err; echo "ref rc=$?"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
