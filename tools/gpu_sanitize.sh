#!/bin/bash
# compute-sanitizer over the GPU parity tests that exercise the cluster / DSMEM / mbarrier / TMA code.
# NOTE (round 2): on this GPU pool compute-sanitizer is closed by the operators ("runs under it have left GPUs needing a
# reset"; gpurun_out/sanitize_*.log hold only that message, rc 86).  Concurrency evidence is therefore the stress test in
# tools/tp_check.py --stress (102 400 exchanges, checksums) and the bounded-spin traps in the kernels.
# usage: tools/gpu_sanitize.sh   (one gpurun call; logs -> gpurun_out/sanitize_*.log)
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
: > gpurun_out/sanitize_summary.txt
one() { # tool, tag, timeout, pytest args...
  tool=$1; tag=$2; to=$3; shift 3
  t0=$(date +%s)
  timeout $to $CS --tool $tool --target-processes all --error-exitcode 9 --launch-timeout 0 \
      python -m pytest "$@" -q -m gpu -x --timeout 0 -p no:cacheprovider > gpurun_out/sanitize_${tool}_${tag}.log 2>&1
  rc=$?
  errs=$(grep -c "========= \(Invalid\|Race\|Error\|Hazard\|Warning: Race\)" gpurun_out/sanitize_${tool}_${tag}.log)
  echo "$tool $tag rc=$rc error_lines=$errs $(( $(date +%s) - t0 ))s :: $(grep -E 'passed|failed|ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/sanitize_${tool}_${tag}.log | tail -2 | tr '\n' ' ')" | tee -a gpurun_out/sanitize_summary.txt
}
one memcheck kv 500 tests/test_gpu_kv.py
one memcheck attn 600 tests/test_gpu_attn.py
one memcheck gemm 900 tests/test_gpu_gemm.py
one memcheck glue 300 tests/test_gpu_glue.py tests/test_gpu_sampling.py
one racecheck kv 400 tests/test_gpu_kv.py
one racecheck attn 600 tests/test_gpu_attn.py
one racecheck gemm 900 tests/test_gpu_gemm.py -k "not streamk"
cat gpurun_out/sanitize_summary.txt
