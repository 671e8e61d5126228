#!/bin/bash
# round-2 evidence run: full GPU suite, full bench line, launch list, ncu --set full of the new kernels
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r2_gpu_tests_v5.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_gpu_tests_v5.log; tail -2 gpurun_out/r2_gpu_tests_v5.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_v3.json 2> gpurun_out/r2_bench_v3.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_v1.json 2> gpurun_out/r2_bench_ref_v1.err; echo "ref rc=$?"
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-others"
$B > gpurun_out/plain_pfb.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench.csv $B > gpurun_out/ncu_launches.log 2>&1
for w in pfbsynth firfff fbank; do
  C="python bench.py --workload $w --steps 1 --warmup 3 --no-cpu --no-e2e --no-others"
  $C > gpurun_out/plain_$w.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'pfb_synth_kernel|fir_fff_kernel|filterbank_kernel' -s 3 -c 1 -o gpurun_out/r2_$w $C > gpurun_out/ncu_$w.log 2>&1
done
ls -la gpurun_out | tail -12
