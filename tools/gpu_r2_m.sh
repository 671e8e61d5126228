#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -k "xlat or xlating or fir or halo or staging" 2>&1 | tail -5
for w in xlat firfff; do python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-others --workload $w 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['config']['workload'][:40], d['value'], d['roofline']['frac'], d.get('parity',{}).get('ok'))"; done
