#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -x -q -k "three_group or fir_fff or xlating or fir_direct" 2>&1 | tail -3
timeout 400 python tools/route_check.py D=16 D=32 2>&1 | grep "ntaps= *\(8\|32\|128\|256\|512\)"
