#!/bin/bash
# TP=4 check: exchange (pull + NVLS) stress, model parity, headline bench at 4 GPUs
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
mkdir -p gpurun_out
timeout 300 $TR --master-port 29541 tools/tp_check.py --stress 2>&1 | grep -v "Warn\|OMP\|\*\*\*" | grep . | tee gpurun_out/tp_check_tp4.txt | tail -12
timeout 400 $TR --master-port 29544 bench.py --gpus 4 --steps 32 --warmup 5 --e2e-steps 16 > gpurun_out/bench_8b_fp8_tp4.json 2> gpurun_out/e4.err; echo rc=$?
python - <<PY
import json
for line in open("gpurun_out/bench_8b_fp8_tp4.json"):
    if line.startswith("{"):
        d=json.loads(line)
        if "value" in d: print(round(d["value"]), d["ms_per_step"], d["e2e"]["value"], [(p["layers_checked"], max(p["logit_rms_err"]), p["bar_frac_of_std"]) for p in d["parity"]["passes"]])
        else: print(d)
PY
grep -v "Warn\|OMP\|\*\*\*" gpurun_out/e4.err | tail -3
