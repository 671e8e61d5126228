#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -x -q -k "three_group or fir_fff or xlating or fir_direct or resampler" 2>&1 | tail -3
timeout 500 python tools/route_check.py D=6 D=10 D=20 D=24 D=12 D=40 2>&1 | grep "ntaps= *\(16\|64\|128\|256\|512\)" | grep -v fff
