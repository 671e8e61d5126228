#!/bin/bash
# round-2 evidence after the fft_filter rewrite: full bench line, reference arm, launch list, ncu --set full of upols2
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_v5.json 2> gpurun_out/r2_bench_v5.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_v3.json 2> gpurun_out/r2_bench_ref_v3.err; echo "ref rc=$?"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e"
$B > gpurun_out/plain_all.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_bench_v2.csv $B > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
bash tools/gpu_ncu_one.sh fftfilt fftfilt_upols r2_upols2
