#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -x -q -k "three_group or fir or xlating or resampler or fftfilt or fft_filter" 2>&1 | tail -3
for w in xlat; do python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-others --workload $w 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['config']['workload'][:40], d['value'], d['roofline']['frac'], (d.get('parity') or {}).get('ok'))"; done
timeout 500 python tools/route_check.py D=8 D=10 D=16 D=20 D=25 2>&1 | grep "not the best\|ccc D=[0-9]* ntaps= *128"
