mkdir -p gpurun_out
for w in fir fftfilt firdec fftvcc; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:'fir_|fftfilt|fft_vcc' -s 3 -c 1 -o gpurun_out/full_$w python bench.py --workload $w --steps 1 --warmup 3 --no-cpu --no-e2e --no-others > gpurun_out/ncu_full_$w.log 2>&1
done
ls -la gpurun_out
