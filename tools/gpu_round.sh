#!/bin/bash
# One gpurun call: GPU parity tests, smoke, the headline bench, then (only if the bench exited 0) the ncu launch list.
# usage: tools/gpu_round.sh [tests] [micro] [bench] [launches] [ncufull]
mkdir -p gpurun_out
: > gpurun_out/summary.txt
want() { [[ " $* " == *" $1 "* ]] ; }
ARGS=" $* "
if [[ "$ARGS" == *" tests "* ]]; then
  for f in test_gpu_kv test_gpu_glue test_gpu_attn test_gpu_gemm test_gpu_sampling test_gpu_model test_gpu_vllm_engine; do
    timeout 600 python -m pytest tests/$f.py -q -m gpu --timeout 300 -s > gpurun_out/$f.log 2>&1
    echo "$f exit=$?" | tee -a gpurun_out/summary.txt
    tail -15 gpurun_out/$f.log
  done
  timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1
  echo "smoke exit=$?" | tee -a gpurun_out/summary.txt
  tail -2 gpurun_out/smoke.log
fi
if [[ "$ARGS" == *" micro "* ]]; then
  timeout 600 python tools/microbench.py ${MICRO_WHAT:-all} --json gpurun_out/microbench.json > gpurun_out/microbench.log 2>&1
  echo "micro exit=$?" | tee -a gpurun_out/summary.txt
  tail -30 gpurun_out/microbench.log
fi
BENCH="python bench.py --steps ${BENCH_STEPS:-64} --warmup 5 ${BENCH_EXTRA}"
if [[ "$ARGS" == *" bench "* ]]; then
  timeout 900 $BENCH > gpurun_out/bench.json 2> gpurun_out/bench.err
  rc=$?
  echo "bench exit=$rc" | tee -a gpurun_out/summary.txt
  tail -c 3000 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
fi
SHORT="python bench.py --steps 3 --warmup 3 --e2e-steps 2 --no-cpu-baseline"
if [[ "$ARGS" == *" launches "* ]]; then
  timeout 600 $SHORT > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:b200:: -s ${LAUNCH_SKIP:-454} -c ${LAUNCH_COUNT:-1362} --csv \
      --log-file gpurun_out/launches.csv $SHORT > gpurun_out/ncu_launches.log 2>&1
  echo "launches exit=$?" | tee -a gpurun_out/summary.txt
fi
if [[ "$ARGS" == *" ncufull "* ]]; then
  timeout 600 $SHORT > gpurun_out/plain2.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:${NCU_KERNEL:-paged_attn_decode_kernel} -s ${NCU_SKIP:-40} -c 3 \
      -f -o gpurun_out/prof_${NCU_TAG:-attn} $SHORT > gpurun_out/ncu_full.log 2>&1
  echo "ncufull exit=$?" | tee -a gpurun_out/summary.txt
  tail -3 gpurun_out/ncu_full.log
fi
cat gpurun_out/summary.txt
