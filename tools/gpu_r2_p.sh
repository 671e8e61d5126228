#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/resamp_check.py 2>&1 | tail -25
for w in resamp interp; do python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-others --workload $w 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['config']['workload'][:40], d['value'], d['roofline']['frac'], (d.get('parity') or {}).get('ok'))"; done
