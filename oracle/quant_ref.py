"""Oracle (test infrastructure): symmetric per-channel weight quantisation and the W8A8 /
weight-only GEMMs.

GEMM arithmetic pinned through llama_ref to HF transformers (+ fake-quant hooks); the quantisation RULE itself has no
external reference (see oracle/__init__.py).  The reference only forwards config keys
(`quantized`, `quantization_type`, `quantization_dtype`; LOADER:885-898) to NxDI.
Restated from the published symmetric scheme:
  scale[n] = amax_k |W[n,k]| / qmax     qmax = 127 (int8) | 448 (OCP e4m3fn)
  Wq = round_half_even(W / scale) clamped (int8)  |  cast_rne_satfinite(W / scale) (fp8)
  activations: same rule per token (row) -- dynamic per-token scale
  out[m,n] = (sum_k Aq[m,k] Wq[n,k]) * a_scale[m] * w_scale[n] (+ bias[n]) -> bf16
Known divergence (documented in DESIGN.md): Neuron's native fp8-e4m3 saturates at 240,
B200 uses OCP e4m3fn (448).
"""
from __future__ import annotations

import torch

FP8_MAX = 448.0
INT8_MAX = 127.0


def _quant_rows(x32, kind: str, tensor_rows=None):
    amax = x32.abs().amax(dim=-1)
    if tensor_rows is not None:  # per_tensor_symmetric: each listed row range (one original tensor) shares one scale
        amax = amax.clone()
        for a, b in tensor_rows:
            amax[a:b] = amax[a:b].max()
    qmax = FP8_MAX if kind == "fp8" else INT8_MAX
    scale = torch.div(amax, torch.full_like(amax, qmax)).float()  # tensor/tensor: true IEEE division
    scale = torch.where(amax > 0, scale, torch.ones_like(scale))
    y = x32 / scale.unsqueeze(-1)
    if kind == "fp8":
        q = y.clamp(-FP8_MAX, FP8_MAX).to(torch.float8_e4m3fn)
    elif kind == "int8":
        q = torch.round(y).clamp(-INT8_MAX, INT8_MAX).to(torch.int8)  # torch.round = half-to-even
    else:
        raise ValueError(kind)
    return q, scale


def quantize_weight_per_channel(w, kind: str, tensor_rows=None):
    """w [N, K] (bf16/fp32) -> (Wq [N,K] fp8|int8, scale [N] fp32).  tensor_rows: 'per_tensor_symmetric' (the
    reference's default quantization_type, LOADER:893-895) -- one scale per listed row range, kept as scale[n]."""
    return _quant_rows(w.float(), kind, tensor_rows)


def quantize_act_per_token(x, kind: str):
    """x [M, K] bf16 -> (Xq [M,K], scale [M] fp32); dynamic symmetric per-token."""
    return _quant_rows(x.float(), kind)


def _f64(w):
    """Weights may be handed over already widened (RefLinear.prepare_f64: the timed CPU baseline converts once,
    outside the timed region, like any CPU runtime would keep its weights in the compute dtype)."""
    return w if w.dtype == torch.float64 else w.to(torch.float64)


def qgemm_ref(xq, x_scale, wq, w_scale, bias=None, residual=None, out_dtype=torch.bfloat16):
    """W8A8: exact integer / fp8 products accumulated in fp64, scaled in fp32."""
    acc = (xq.to(torch.float64) @ _f64(wq).t()).float()
    y = acc * x_scale.view(-1, 1) * w_scale.view(1, -1)
    if bias is not None:
        y = y + bias.float().view(1, -1)
    if residual is not None:
        y = y + residual.float()
    return y.to(out_dtype)


def gemm_ref(x, w, bias=None, residual=None, out_dtype=torch.bfloat16):
    """Plain bf16 GEMM: out = x @ w^T (+bias)(+residual), fp32+ accumulate, one rounding."""
    y = (x.to(torch.float64) @ _f64(w).t()).float()
    if bias is not None:
        y = y + bias.float().view(1, -1)
    if residual is not None:
        y = y + residual.float()
    return y.to(out_dtype)


def wo_gemm_ref(x, wq, w_scale, bias=None, residual=None, out_dtype=torch.bfloat16):
    """Weight-only int8: bf16 activations x (int8 weights dequantised per channel)."""
    acc = (x.to(torch.float64) @ _f64(wq).t()).float()
    y = acc * w_scale.view(1, -1)
    if bias is not None:
        y = y + bias.float().view(1, -1)
    if residual is not None:
        y = y + residual.float()
    return y.to(out_dtype)
