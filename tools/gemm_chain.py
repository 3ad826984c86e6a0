#!/usr/bin/env python
"""Cost of a decode GEMM inside a dependent chain: [per-token quantise -> GEMM] x R with R distinct weights (so every
weight byte comes from HBM), captured in a CUDA graph and replayed; reports us per (quantise + GEMM) pair for each
split_k.  This is the number that matters in the decode step (kernel boundaries and the reduction tail included), unlike a
back-to-back microbenchmark of the GEMM alone.

    python tools/gemm_chain.py M N K [split_k ...]      (env knobs of the library apply, e.g. B200_GEMM_FORCE_OCC=1)
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_neuron_b200 import ops  # noqa: E402


def main():
    m, n, k = (int(x) for x in sys.argv[1:4])
    splits = [int(x) for x in sys.argv[4:]] or [0]
    ops.device_check(0)
    reps = max(8, int(600e6 // (n * k)))  # > 4x L2 of distinct weights
    ws = [torch.randn(n, k, device="cuda", dtype=torch.bfloat16).to(torch.float8_e4m3fn) for _ in range(reps)]
    x = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    sw = torch.ones(n, device="cuda")
    floor_us = (n * k + m * k + 2 * m * n) / 6.46e6
    for sk in splits:
        try:
            plan = ops.gemm_decode_plan(m, n, k, 1, split_k=sk)
        except Exception as e:  # noqa: BLE001
            print(f"split_k={sk}: {e}")
            continue

        def chain():
            for w in ws:
                aq, a_s = ops.quantize_per_token(x, "fp8")
                ops.gemm(aq, w, a_s, sw, out=out, split_k=sk)

        def only_q():
            for _ in ws:
                ops.quantize_per_token(x, "fp8")

        res = []
        for body in (chain, only_q):
            body()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                body()
            for _ in range(3):
                g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            res.append(e0.elapsed_time(e1) * 1e3 / (10 * reps))
        pair, q = res
        print(f"M={m} N={n} K={k} split_k={sk} plan={plan}: pair {pair:6.2f} us, quantise alone {q:5.2f} us, "
              f"GEMM adds {pair - q:6.2f} us (HBM floor {floor_us:5.2f} us)", flush=True)


if __name__ == "__main__":
    main()
