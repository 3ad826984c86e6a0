#!/usr/bin/env python
"""Kernel-level micro-benchmarks (CUDA events, inputs rotated through > L2-size working sets).

    python tools/microbench.py attn|gemm|all [--json out.json]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_neuron_b200 import ops  # noqa: E402

PEAKS = {}
try:
    PEAKS = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
except Exception:
    pass
HBM = PEAKS.get("hbm_gbs", 6650.0)
# dense FP8 peak measured by tools/fp8_peak.py on this pool (cuBLASLt e4m3, sustained); None until that file is committed
FP8_PEAK = None
try:
    FP8_PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles",
                                           "fp8_peak.json"))).get("fp8_tflops_sustained")
except Exception:
    pass


def time_calls(fns, iters=3, warmup=1, graph=True):
    """fns: list of closures run back to back (one 'round'); returns avg ms per closure.  With graph=True the round is
    captured into a CUDA graph and replayed, so the number is device time, not Python/ctypes launch rate."""
    for _ in range(warmup):
        for f in fns:
            f()
    torch.cuda.synchronize()
    if graph:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for f in fns:
                f()
        run = g.replay
        run()
        torch.cuda.synchronize()
    else:
        def run():
            for f in fns:
                f()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (iters * len(fns))


def bench_attn(b=64, hq=32, hkv=8, d=128, ctx=2048, layers=8, quotas=(0, 8, 16)):
    bs = 32
    nblk = ctx // bs
    nb = b * nblk + 1
    g = torch.Generator(device="cuda").manual_seed(11)
    pools = [
        (torch.randn(nb, hkv, bs, d, generator=g, device="cuda", dtype=torch.bfloat16),
         torch.randn(nb, hkv, bs, d, generator=g, device="cuda", dtype=torch.bfloat16))
        for _ in range(layers)
    ]
    perm = torch.randperm(nb - 1, generator=torch.Generator().manual_seed(7)) + 1
    bt = perm[: b * nblk].reshape(b, nblk).to(torch.int32).cuda()
    ctxl = torch.full((b,), ctx, dtype=torch.int32, device="cuda")
    q = torch.randn(b, hq, d, generator=g, device="cuda", dtype=torch.bfloat16)
    out = torch.empty_like(q)
    ws = torch.empty(ops.paged_attn_decode_workspace_bytes(b, hq, hkv, d, nblk, 0), dtype=torch.uint8, device="cuda")
    alg = b * ctx * 2 * hkv * d * 2 + 2 * b * hq * d * 2 + b * nblk * 4
    res = []
    for cb in quotas:
        fns = [
            (lambda kp=kp, vp=vp: ops.paged_attn_decode(q, kp, vp, bt, ctxl, d**-0.5, out=out, workspace=ws, min_quota=cb))
            for kp, vp in pools
        ]
        ms = time_calls(fns)
        gbs = alg / ms / 1e6
        res.append({"kernel": "paged_attn_decode", "B": b, "Hq": hq, "Hkv": hkv, "D": d, "ctx": ctx, "min_quota": cb,
                    "ms": ms, "GBps": gbs, "frac_hbm": gbs / HBM})
        print(res[-1], flush=True)
    return res


def bench_gemm(shapes=None, kinds=("fp8", "bf16"), pdl=True, splits=(0,)):
    shapes = shapes or [(64, 6144, 4096), (64, 4096, 4096), (64, 28672, 4096), (64, 4096, 14336), (64, 128256, 4096)]
    ops.set_pdl(pdl)
    res = []
    for kind in kinds:
        for m, n, k in shapes:
            eb = 2 if kind == "bf16" else 1
            dtc = {"bf16": 0, "fp8": 1, "int8wo": 4}[kind]
            copies = max(2, int(300e6 // (n * k * eb)) + 1)
            if kind == "bf16":
                ws_ = [torch.randn(n, k, device="cuda", dtype=torch.bfloat16) for _ in range(copies)]
                a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
                sa = sw = None
            elif kind == "int8wo":
                ws_ = [torch.randint(-127, 128, (n, k), device="cuda", dtype=torch.int8) for _ in range(copies)]
                a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
                sa, sw = None, torch.ones(n, device="cuda")
            else:
                ws_ = [torch.randn(n, k, device="cuda", dtype=torch.bfloat16).to(torch.float8_e4m3fn) for _ in range(copies)]
                a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16).to(torch.float8_e4m3fn)
                sa = torch.ones(m, device="cuda")
                sw = torch.ones(n, device="cuda")
            out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
            nbytes = ops.gemm_workspace_bytes(m, n, k, dtc, 0)
            wsp = torch.zeros(max(nbytes, 16), dtype=torch.uint8, device="cuda")  # stream-K counters start at zero
            for sk in splits:
                try:
                    plan = ops.gemm_decode_plan(m, n, k, dtc, split_k=sk)
                except Exception:
                    continue
                fns = [(lambda w=w: ops.gemm(a, w, sa, sw, out=out, workspace=wsp, split_k=sk)) for w in ws_]
                ms = time_calls(fns, iters=5)
                alg = n * k * eb + n * 4 + m * k * (2 if kind == "int8wo" else eb) + m * 4 + 2 * m * n
                gbs = alg / ms / 1e6
                res.append({"kernel": f"gemm_{kind}", "pdl": pdl, "M": m, "N": n, "K": k, "split_k": sk, "us": round(ms * 1e3, 2),
                            "GBps": round(gbs, 1), "frac_hbm": round(gbs / HBM, 4),
                            "TFLOPs": round(2 * m * n * k / ms / 1e9, 1), "plan": plan})
                print(res[-1], flush=True)
            del ws_
    return res


def bench_configs():
    """Kernel-level numbers at the PER-GPU shard shapes of BASELINE configs 4 and 5 (one GPU stands in for one TP rank;
    no collective is timed):
      cfg5: Llama-3.3-70B FP8, TP=8 -> Hq=8, Hkv=1 per rank, B=64, context 2k..32k (flash-decoding split over the context);
      cfg4: Qwen2.5-32B INT8 weight-only, TP=8, B=128 -> Hq=5, Hkv=1; linears 5120 x {896, 640->5120, 2x3456, 3456->5120}."""
    res = []
    for ctx in (2048, 4096, 8192, 16384, 32768):
        layers = max(2, min(8, (1 << 30) // (64 * ctx * 2 * 128 * 2)))  # > 4x L2 of distinct KV per round
        r = bench_attn(b=64, hq=8, hkv=1, d=128, ctx=ctx, layers=layers, quotas=(0,))
        for x in r:
            x["config"] = "cfg5 llama-3.3-70b tp8 shard"
        res += r
    r = bench_attn(b=128, hq=5, hkv=1, d=128, ctx=2048, layers=8, quotas=(0,))
    for x in r:
        x["config"] = "cfg4 qwen2.5-32b tp8 shard"
    res += r
    # cfg4 linears, weight-only int8 (bf16 activations), M = 128
    m = 128
    for n, k, name in ((896, 5120, "qkv"), (5120, 640, "o"), (6912, 5120, "gate_up"), (5120, 3456, "down"), (19008, 5120, "lm_head")):
        copies = max(4, int(300e6 // (n * k)) + 1)
        ws_ = [torch.randint(-127, 128, (n, k), device="cuda", dtype=torch.int8) for _ in range(copies)]
        a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
        sw = torch.rand(n, device="cuda") * 1e-3
        out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        fns = [(lambda w=w: ops.gemm(a, w, None, sw, out=out)) for w in ws_]
        ms = time_calls(fns, iters=5)
        alg = n * k + n * 4 + m * k * 2 + 2 * m * n
        gbs = alg / ms / 1e6
        res.append({"kernel": "gemm_int8wo", "config": "cfg4 qwen2.5-32b tp8 shard", "linear": name, "M": m, "N": n, "K": k,
                    "us": round(ms * 1e3, 2), "GBps": round(gbs, 1), "frac_hbm": round(gbs / HBM, 4),
                    "plan": ops.gemm_decode_plan(m, n, k, 4)})
        print(res[-1], flush=True)
        del ws_
    return res


def bench_prefill(attn_only=False):
    """Prefill-side kernels of Llama-3.1-8B: gemm_streamk_kernel at M = 512 / 2048 / 8192 for
    qkv / gate_up / down in bf16 and fp8 (TFLOP/s against the measured sustained bf16 cuBLAS figure), and
    paged_attn_prefill_kernel at S_new + cached = 2048 (0 and 1024 cached tokens) and 8192."""
    tf_peak = PEAKS.get("bf16_tflops_sustained", 1432.9)
    res = []
    for kind in (() if attn_only else ("bf16", "fp8")):
        for m in (512, 2048, 8192):
            for n, k, name in ((6144, 4096, "qkv"), (28672, 4096, "gate_up"), (4096, 14336, "down")):
                if kind == "bf16":
                    w = torch.randn(n, k, device="cuda", dtype=torch.bfloat16)
                    a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
                    sa = sw = None
                else:
                    w = torch.randn(n, k, device="cuda", dtype=torch.bfloat16).to(torch.float8_e4m3fn)
                    a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16).to(torch.float8_e4m3fn)
                    sa, sw = torch.ones(m, device="cuda"), torch.ones(n, device="cuda")
                out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
                dtc = {"bf16": 0, "fp8": 1}[kind]
                wsp = torch.zeros(max(ops.gemm_workspace_bytes(m, n, k, dtc, 0), 16), dtype=torch.uint8, device="cuda")
                fns = [lambda: ops.gemm(a, w, sa, sw, out=out, workspace=wsp)] * 4
                ms = time_calls(fns, iters=5)
                tf = 2 * m * n * k / ms / 1e9
                res.append({"kernel": f"gemm_streamk_{kind}", "linear": name, "M": m, "N": n, "K": k, "us": round(ms * 1e3, 1),
                            "TFLOPs": round(tf, 1), "frac_bf16_sustained_peak": round(tf / tf_peak, 3)})
                if kind == "fp8" and FP8_PEAK:  # an FP8 GEMM is judged against the FP8 tensor peak, not the bf16 one
                    res[-1]["frac_fp8_sustained_peak"] = round(tf / FP8_PEAK, 3)
                print(res[-1], flush=True)
    hq, hkv, d, bs = 32, 8, 128, 32
    cases = ((2048, 0), (2048, 1024), (8192, 0), (8192, 4096), (32768, 0))
    if os.environ.get("PREFILL_CASES"):  # e.g. "8192:0" -- a single shape under ncu
        cases = tuple(tuple(int(x) for x in c.split(":")) for c in os.environ["PREFILL_CASES"].split(","))
    for total, cached in cases:
        s_new = total - cached
        nblk = total // bs
        g = torch.Generator(device="cuda").manual_seed(3)
        kp = torch.randn(nblk + 1, hkv, bs, d, generator=g, device="cuda", dtype=torch.bfloat16)
        vp = torch.randn(nblk + 1, hkv, bs, d, generator=g, device="cuda", dtype=torch.bfloat16)
        bt = (torch.randperm(nblk, generator=torch.Generator().manual_seed(1)) + 1).to(torch.int32).cuda()
        q = torch.randn(s_new, hq, d, generator=g, device="cuda", dtype=torch.bfloat16)
        out = torch.empty_like(q)
        fns = [lambda: ops.paged_attn_prefill(q, kp, vp, bt, cached, d ** -0.5, out=out)] * 4
        ms = time_calls(fns, iters=5)
        # causal flops: every new token attends to cached + its own prefix: QK^T and PV, 2 flops per MAC
        pairs = s_new * cached + s_new * (s_new + 1) // 2
        tf = 4 * pairs * hq * d / ms / 1e9
        res.append({"kernel": "paged_attn_prefill", "S_new": s_new, "cached": cached, "Hq": hq, "Hkv": hkv, "D": d,
                    "us": round(ms * 1e3, 1), "TFLOPs_causal": round(tf, 1), "frac_bf16_sustained_peak": round(tf / tf_peak, 3)})
        print(res[-1], flush=True)
    return res


def bench_ttft(quant="fp8", prompts=(512, 2048, 8192)):
    """Time to first token of one request through the model's prefill path (embedding -> 32 layers with the stream-K GEMMs,
    fused RoPE / KV write, tcgen05 prefill attention -> last-row lm_head -> argmax), Llama-3.1-8B, random-init weights."""
    from vllm_neuron_b200.config import PRESETS
    from vllm_neuron_b200.model import B200CausalLM

    shape = PRESETS["llama-3.1-8b"]
    mml = max(prompts) + 32
    mb = -(-mml // 32)
    nb = mb + 1
    model = B200CausalLM(shape, None, quant=quant, num_blocks=nb, max_num_seqs=1, max_model_len=mml, device="cuda", seed=0)
    model.block_table[0, :mb] = torch.arange(1, mb + 1, dtype=torch.int32, device="cuda")
    res = []
    for n in prompts:
        toks = torch.randint(0, shape.vocab, (n,), generator=torch.Generator().manual_seed(n)).cuda()
        for cached in (0, n // 2):
            new = toks[cached:]
            model.prefill(0, new, n, cached)  # warm-up (plans, tensor maps)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 3
            e0.record()
            for _ in range(reps):
                tok = model.prefill(0, new, n, cached)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            # dense flops of the new tokens: 2 x params(layers + head row) x tokens + attention
            qd, kd = shape.num_q_heads * shape.head_dim, shape.num_kv_heads * shape.head_dim
            per_tok = 2 * shape.layers * ((qd + 2 * kd) * shape.hidden + shape.hidden * qd + 3 * shape.intermediate * shape.hidden)
            s_new = n - cached
            attn = 4 * (s_new * cached + s_new * (s_new + 1) // 2) * shape.num_q_heads * shape.head_dim * shape.layers
            tf = (per_tok * s_new + attn) / ms / 1e9
            res.append({"kernel": "ttft_prefill", "model": shape.name, "quant": quant, "prompt": n, "cached_prefix": cached,
                        "ms": round(ms, 3), "prefill_tokens_per_s": round(s_new / ms * 1e3), "TFLOPs": round(tf, 1),
                        "frac_bf16_sustained_peak": round(tf / PEAKS.get("bf16_tflops_sustained", 1432.9), 3),
                        "first_token": int(tok)})
            print(res[-1], flush=True)
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("what", nargs="?", default="all")
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    ops.device_check(0)
    out = []
    if a.what in ("attn", "all"):
        out += bench_attn()
    if a.what in ("gemm", "all"):
        out += bench_gemm()
        out += bench_gemm(kinds=("fp8",), pdl=False)
        ops.set_pdl(True)
    if a.what == "gemmcmp":
        out += bench_gemm(shapes=[(64, 4096, 4096), (128, 4096, 4096), (64, 28672, 4096), (128, 28672, 4096), (128, 6912, 5120),
                                  (128, 5120, 640)], kinds=("fp8", "bf16", "int8wo"))
    if a.what == "configs":
        out += bench_configs()
    if a.what == "ttft":
        out += bench_ttft()
    if a.what == "prefillattn":
        out += bench_prefill(attn_only=True)
    if a.what == "prefill":
        out += bench_prefill()
    if a.what == "gemmsweep":
        out += bench_gemm(kinds=("fp8",), splits=(0, 1, 2, 3, 4, 5, 6, 7, 8))
    if a.json:
        json.dump(out, open(a.json, "w"), indent=1)
