#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -x -q -k "three_group or fir or xlating or resampler or fftfilt or fft_filter" 2>&1 | tail -3
timeout 500 python tools/route_check.py D=3 D=5 D=25 2>&1 | grep "ntaps= *\(64\|128\|512\)"
