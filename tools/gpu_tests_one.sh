#!/bin/bash
# run selected GPU tests: $1 = log name, rest = pytest args
mkdir -p gpurun_out
log=gpurun_out/$1; shift
python -m pytest "$@" -q -s -m gpu > $log 2>&1; echo "pytest rc=$?" >> $log
grep -E "passed|failed|rc=" $log | tail -5
