#!/bin/bash
# GPU parity tests, each file under its own timeout so a hang cannot eat the box.
mkdir -p gpurun_out
: > gpurun_out/summary.txt
for f in "$@"; do
  timeout 420 python -m pytest tests/$f.py -q -m gpu --timeout 120 > gpurun_out/$f.log 2>&1
  echo "$f exit=$?" | tee -a gpurun_out/summary.txt
  tail -25 gpurun_out/$f.log
done
