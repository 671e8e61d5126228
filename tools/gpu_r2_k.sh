#!/bin/bash
# three-group plain overlap-save kernel: parity tests of every block that can reach it, then the routes old vs new
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -k "fftfilt or fft_filter or fir or staging or xlat" 2>&1 | tail -5
timeout 600 python tools/route_check.py > gpurun_out/route_check_v2.txt 2>&1; grep -c "not the best" gpurun_out/route_check_v2.txt
timeout 300 python tools/fftfilt_check.py 1399 1800 2048 2049 2>&1 | grep "upols=0\|2049"
