#!/usr/bin/env python
"""Decode attention at TP-shard head counts: time vs the minimum blocks-per-worker quota of the flash-decoding split."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import importlib.util

spec = importlib.util.spec_from_file_location("mb", os.path.join(os.path.dirname(os.path.abspath(__file__)), "microbench.py"))
mb = importlib.util.module_from_spec(spec)
sys.argv = ["x"]
spec.loader.exec_module(mb)
ops = mb.ops
d, bs, ctx, layers = 128, 32, 2048, 8
for (hq, hkv, b) in ((4, 1, 64), (8, 2, 64), (16, 4, 64), (8, 1, 64), (5, 1, 128)):
    nblk = ctx // bs
    nb = b * nblk + 1
    g = torch.Generator(device="cuda").manual_seed(11)
    pools = [(torch.randn(nb, hkv, bs, d, generator=g, device="cuda", dtype=torch.bfloat16),
              torch.randn(nb, hkv, bs, d, generator=g, device="cuda", dtype=torch.bfloat16)) for _ in range(layers)]
    bt = (torch.randperm(nb - 1, generator=torch.Generator().manual_seed(7)) + 1)[: b * nblk].reshape(b, nblk).to(torch.int32).cuda()
    ctxl = torch.full((b,), ctx, dtype=torch.int32, device="cuda")
    q = torch.randn(b, hq, d, generator=g, device="cuda", dtype=torch.bfloat16)
    out = torch.empty_like(q)
    alg = b * ctx * 2 * hkv * d * 2 + 2 * b * hq * d * 2 + b * nblk * 4
    for cb in (1, 2, 3, 4, 6, 8, 12):
        ws = torch.empty(ops.paged_attn_decode_workspace_bytes(b, hq, hkv, d, nblk, cb), dtype=torch.uint8, device="cuda")
        fns = [(lambda kp=kp, vp=vp: ops.paged_attn_decode(q, kp, vp, bt, ctxl, d ** -0.5, out=out, workspace=ws, min_quota=cb))
               for kp, vp in pools]
        ms = mb.time_calls(fns)
        print(f"B {b} Hq {hq} Hkv {hkv} quota {cb}: {ms * 1e3:.1f} us  {alg / ms / 1e6 / mb.HBM:.3f} of HBM", flush=True)
