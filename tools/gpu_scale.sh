#!/bin/bash
# usage: tools/gpu_scale.sh N [extra bench args]   -- one bench.py run at N GPUs under torchrun, bounded
N=$1; shift
mkdir -p gpurun_out
if [ "$N" = "1" ]; then
  timeout 600 python bench.py --gpus 1 "$@" > gpurun_out/bench_tp1.json 2> gpurun_out/bench_tp1.err
else
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N "$@" > gpurun_out/bench_tp$N.json 2> gpurun_out/bench_tp$N.err
fi
echo "rc=$?"
python - <<PY
import json
for line in open("gpurun_out/bench_tp$N.json"):
    line=line.strip()
    if line.startswith("{"):
        d=json.loads(line)
        print({k:d[k] for k in ("value","n_gpus","ms_per_step")}, "e2e", round(d["e2e"]["value"],1), "attn frac", d["roofline"]["frac"], "step frac", d["roofline"]["step"]["frac_of_step_roofline"])
PY
grep -v Warn gpurun_out/bench_tp$N.err | grep -v "^\*\*\*\|OMP_NUM" | tail -5
