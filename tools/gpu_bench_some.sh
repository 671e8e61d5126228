#!/bin/bash
# $1 = tag; rest = workloads: quick device-resident numbers (no cpu, no e2e, no others)
mkdir -p gpurun_out
tag=$1; shift
for w in "$@"; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e --no-others 2>gpurun_out/bench_${tag}_$w.err | python -c "
import sys,json
d=json.loads(sys.stdin.read()); r=d['roofline']
print('%-9s %9.0f Msps  ms/step %.3f  hbm_frac %.3f  fp32_tflops %.1f  parity %s' % ('$w', d['value'], d['ms_per_step'], r['achieved']/r['peak'] if r['unit']=='GB/s' else -1, r['fp32_tflops'], (d.get('parity') or {}).get('ok')))
" | tee -a gpurun_out/bench_${tag}.txt
done
