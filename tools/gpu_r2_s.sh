#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -k "fftfilt or fft_filter or fir or staging or xlat or xlating or resampler" 2>&1 | tail -5
timeout 600 python tools/route_check.py > gpurun_out/route_check_v6.txt 2>&1; grep "D=[248] ntaps= *\(16\|64\|128\|512\)" gpurun_out/route_check_v6.txt
