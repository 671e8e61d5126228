#!/bin/bash
# full GPU suite; log -> gpurun_out/$1
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s > gpurun_out/$1 2>&1; echo "pytest rc=$?" >> gpurun_out/$1
grep -E "passed|failed|rc=" gpurun_out/$1 | tail -5
