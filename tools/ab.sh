#!/bin/bash
# A/B of an environment knob on the headline bench (short runs, no CPU leg): tools/ab.sh VAR v1 v2 ...
mkdir -p gpurun_out
VAR=$1; shift
for v in "$@"; do
  env $VAR=$v timeout 600 python bench.py --steps 32 --warmup 5 --e2e-steps 4 --no-cpu-baseline > gpurun_out/ab_${VAR}_$v.json 2> gpurun_out/ab_${VAR}_$v.err
  python - <<P
import json
try:
    d=json.load(open("gpurun_out/ab_${VAR}_$v.json"))
    print("$VAR=$v", round(d["value"],1), "tok/s", round(d["ms_per_step"],4), "ms; attn", d["roofline"]["mean_us_per_launch"], "us frac", d["roofline"]["frac"])
except Exception as e:
    print("$VAR=$v failed", e); print(open("gpurun_out/ab_${VAR}_$v.err").read()[-1500:])
P
done
