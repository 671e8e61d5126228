#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -k "logpwr or no_alloc or static" 2>&1 | tail -3
for m in 0 1; do GRB200_LOGPWR_FUSED=$m python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-others --workload logpwr 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['config']['workload'][:40], d['value'], d['roofline']['frac'], d.get('parity'))"; done
