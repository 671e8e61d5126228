#!/bin/bash
timeout 900 python -m pytest tests -m gpu -x -q -k "resampler or interp" 2>&1 | tail -4
timeout 300 python tools/resamp_check.py 2>&1 | tail -7
