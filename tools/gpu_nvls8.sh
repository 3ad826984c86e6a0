TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
mkdir -p gpurun_out
timeout 300 $TR --master-port 29541 tools/tp_check.py --stress 2>&1 | grep -v "Warn\|OMP\|\*\*\*" | grep . | tee gpurun_out/tp_check_nvls_tp8.txt | tail -8
timeout 200 $TR --master-port 29542 tools/tp_check.py --bench 2>&1 | grep "exchange bench" | tee gpurun_out/exchange_bench_tp8.txt
timeout 300 $TR --master-port 29543 bench.py --gpus 8 --quant bf16 --steps 16 --warmup 3 --e2e-steps 4 > gpurun_out/bench_8b_bf16_tp8_nvls.json 2> gpurun_out/e1.err; echo rc=$?
timeout 300 $TR --master-port 29544 bench.py --gpus 8 --steps 32 --warmup 5 --e2e-steps 16 > gpurun_out/bench_8b_fp8_tp8_nvls.json 2> gpurun_out/e2.err; echo rc=$?
python - <<PY
import json
for f in ["bench_8b_bf16_tp8_nvls","bench_8b_fp8_tp8_nvls"]:
    for line in open(f"gpurun_out/{f}.json"):
        if line.startswith("{"):
            d=json.loads(line)
            if "value" in d: print(f, round(d["value"]), d["ms_per_step"], [(p["layers_checked"], max(p["logit_rms_err"]), p["bar_frac_of_std"]) for p in d["parity"]["passes"]])
            else: print(d)
PY
