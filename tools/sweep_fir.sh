#!/bin/bash
# FIR plan sweep on the GPU box: GRB200_FIR_PLAN=R,NT,NBUF (NBUF 3 = warp-private pipelines)
mkdir -p gpurun_out
out=gpurun_out/sweep_fir.txt; : > $out
for w in fir firdec; do
  for plan in default 16,128,3 16,256,3 8,128,3 8,256,3 16,256,2 16,128,2 8,256,2 8,128,2 16,256,1 16,128,1 8,256,1 8,128,1 8,64,2; do
    if [ $plan = default ]; then unset GRB200_FIR_PLAN; else export GRB200_FIR_PLAN=$plan; fi
    r=$(timeout 120 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu --no-e2e --no-others 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['roofline']['fp32_tflops'],1))")
    echo "$w $plan $r" >> $out
  done
done
cat $out
