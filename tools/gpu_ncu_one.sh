#!/bin/bash
# $1 = workload, $2 = kernel regex, $3 = output tag: plain run first, then one ncu --set full capture with source
mkdir -p gpurun_out
C="python bench.py --workload $1 --steps 1 --warmup 3 --no-cpu --no-e2e --no-others"
$C > gpurun_out/plain_$3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"$2" -s 3 -c 1 -f -o gpurun_out/$3 $C > gpurun_out/ncu_$3.log 2>&1
echo "rc=$?"; ls -la gpurun_out/$3.ncu-rep
