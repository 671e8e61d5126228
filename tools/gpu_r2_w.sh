#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -x -q -k "pfb_against_oracle" 2>&1 | tail -3
timeout 300 python tools/pfb_m_check.py 64x16 64x40 64x64 32x48 256x33 8x64
