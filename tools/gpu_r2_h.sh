#!/bin/bash
# round-2 final-code evidence: full GPU suite + full bench line + reference arm (no profiler in this call)
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r2_gpu_tests_v6.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_gpu_tests_v6.log; tail -2 gpurun_out/r2_gpu_tests_v6.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_v4.json 2> gpurun_out/r2_bench_v4.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_v2.json 2> gpurun_out/r2_bench_ref_v2.err; echo "ref rc=$?"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
