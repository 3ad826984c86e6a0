#!/bin/bash
# One multi-GPU gpurun call: tools/gpu_multi.sh N [trial]
#   TP exchange stress + model parity at N ranks, headline bench at TP=N, BASELINE configs[3] (Qwen2.5-32B INT8 weight-only,
#   batch 128) and configs[4] (Llama-3.3-70B FP8 context sweep) as whole-model runs.  "trial" shrinks the models (4 layers)
#   so the plumbing can be checked on 2 GPUs.
N=$1; MODE=$2
mkdir -p gpurun_out
: > gpurun_out/multi_summary.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
L=""; SW="2048,4096,8192,16384,32768"; ST=16
if [ "$MODE" = "trial" ]; then L="--layers 4"; SW="2048,4096"; ST=8; fi
run() { # name, timeout, cmd...
  name=$1; to=$2; shift 2
  t0=$(date +%s)
  timeout $to "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  rc=$?
  echo "$name rc=$rc $(( $(date +%s) - t0 ))s" | tee -a gpurun_out/multi_summary.txt
  grep -v "Warn\|^\*\*\*\|OMP_NUM\|^$" gpurun_out/$name.err | tail -4
}
[ -z "$SKIP_STRESS" ] && run tp_check_tp$N 400 $TR --master-port 29541 tools/tp_check.py --stress
[ -z "$SKIP_BF16" ] && run bench_8b_bf16_tp$N 400 $TR --master-port 29545 bench.py --gpus $N --quant bf16 --steps 16 --warmup 3 --e2e-steps 4 $L
run bench_8b_fp8_tp$N 400 $TR --master-port 29542 bench.py --gpus $N --steps 32 --warmup 5 --e2e-steps 16 $L
run bench_cfg4_qwen32b_int8wo_b128_tp$N 500 $TR --master-port 29543 bench.py --gpus $N --model qwen2.5-32b --quant int8_wo --batch 128 --ctx 2048 --steps $ST --warmup 3 --e2e-steps 8 $L
run bench_cfg5_llama70b_fp8_sweep_tp$N 900 $TR --master-port 29544 bench.py --gpus $N --model llama-3.3-70b --quant fp8 --batch 64 --sweep-ctx $SW --steps $ST --warmup 3 --e2e-steps 4 $L
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/bench_*tp$N.json")):
    for line in open(f):
        line = line.strip()
        if not line.startswith("{"): continue
        d = json.loads(line)
        if "value" not in d: print(f, d); continue
        print(f.split("/")[-1], d["config"]["workload"][:40], "ctx", d["config"]["context"], "tok/s", round(d["value"]), "ms", round(d["ms_per_step"], 3),
              "e2e", round(d["e2e"]["value"]), "attn", d["roofline"]["frac"], "step", d["roofline"]["step"]["frac_of_step_roofline"],
              "parity", (d.get("parity") or {}).get("tp_check"), [(p["layers_checked"], max(p["logit_rms_err"] or [0]), p["logit_std"][:1]) for p in (d.get("parity") or {}).get("passes", [])])
PY
cat gpurun_out/multi_summary.txt
