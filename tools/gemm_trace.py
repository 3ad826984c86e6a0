#!/usr/bin/env python
"""Phase timeline of the decode GEMM from in-kernel %globaltimer stamps (b200_gemm_set_trace).

    python tools/gemm_trace.py [M N K [split_k]]

Launches a PDL chain [quantize-like producer kernel -> GEMM] x R and prints, for the LAST GEMM of the chain, when each
phase happened relative to the moment the dependency resolved (median over CTAs; min..max).  Diagnostic only.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_neuron_b200 import _cabi, ops  # noqa: E402

NAMES = ["cta start", "dependency resolved", "first stage landed", "last MMA retired", "partials pushed",
         "cluster barrier passed", "outputs stored"]


def main():
    m, n, k = (int(x) for x in sys.argv[1:4]) if len(sys.argv) >= 4 else (64, 4096, 4096)
    sk = int(sys.argv[4]) if len(sys.argv) >= 5 else 0
    ops.device_check(0)
    lib = _cabi.load()
    plan = ops.gemm_decode_plan(m, n, k, 1, split_k=sk)
    ctas = -(-n // 128) * plan["split"]
    reps = 6
    ws = [torch.randn(n, k, device="cuda", dtype=torch.bfloat16).to(torch.float8_e4m3fn) for _ in range(reps)]
    x = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    sw = torch.ones(n, device="cuda")
    trace = torch.zeros(ctas * 24, dtype=torch.int64, device="cuda")

    def chain():
        for w in ws:
            aq, a_s = ops.quantize_per_token(x, "fp8")  # a small producer kernel in front of every GEMM
            ops.gemm(aq, w, a_s, sw, out=out, split_k=sk)

    chain()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    lib.b200_gemm_set_trace(trace.data_ptr())
    with torch.cuda.graph(g):
        chain()
    lib.b200_gemm_set_trace(None)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    kb = trace[ctas * 8:].view(ctas, 16).cpu().double()
    t = trace[: ctas * 8].view(ctas, 8).cpu().double()
    t0 = t[:, 1].median()  # dependency resolved
    print(f"M={m} N={n} K={k} plan={plan} ctas={ctas}")
    for i, name in enumerate(NAMES):
        col = t[:, i]
        col = col[col > 0]
        if col.numel() == 0:
            continue
        r = (col - t0) / 1e3
        print(f"  {name:26s} median {r.median():7.2f} us   min {r.min():7.2f}   max {r.max():7.2f}")
    for i in range(8):
        w_, a_ = kb[:, i], kb[:, 8 + i]
        if (w_ > 0).any():
            print(f"  k-block {i}: W seen {((w_[w_ > 0] - t0) / 1e3).median():6.2f} us, A seen {((a_[a_ > 0] - t0) / 1e3).median():6.2f} us (medians)")


if __name__ == "__main__":
    main()
