for occ in 0 1; do
  echo "== FORCE_OCC=$occ"
  B200_GEMM_FORCE_OCC=$occ python tools/gemm_chain.py 64 6144 4096 0 2 3 4 5 6 2>&1 | grep -v Warn
  B200_GEMM_FORCE_OCC=$occ python tools/gemm_chain.py 64 4096 4096 0 2 3 4 5 6 8 2>&1 | grep -v Warn
  B200_GEMM_FORCE_OCC=$occ python tools/gemm_chain.py 64 4096 14336 0 3 4 5 6 7 2>&1 | grep -v Warn
  B200_GEMM_FORCE_OCC=$occ python tools/gemm_chain.py 64 28672 4096 0 1 2 2>&1 | grep -v Warn
done
