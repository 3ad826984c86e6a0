#!/bin/bash
# One gpurun call: GPU parity tests (each file under its own timeout so a hang cannot eat the box), then micro-benchmarks.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
for f in test_gpu_kv test_gpu_glue test_gpu_attn test_gpu_gemm; do
  timeout 420 python -m pytest tests/$f.py -q -m gpu -x --timeout 120 > gpurun_out/$f.log 2>&1
  echo "$f exit=$?" | tee -a gpurun_out/summary.txt
  tail -5 gpurun_out/$f.log
done
timeout 300 python tools/microbench.py all --json gpurun_out/microbench.json > gpurun_out/microbench.log 2>&1
echo "microbench exit=$?" | tee -a gpurun_out/summary.txt
tail -30 gpurun_out/microbench.log
