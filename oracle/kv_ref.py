"""Oracle (test infrastructure): block-KV cache write, RoPE, and small glue ops.

Arithmetic (RoPE, RMSNorm, SiLU) pinned through llama_ref to HF transformers (see oracle/__init__.py): the reference's KV
write happens inside NxDI's compiled graph.  What the reference fixes -- and what
this file restates -- is the *indexing contract* of its host side:
  * slot = block_id * block_size + offset   (RUNNER:757-763, 819-823)
  * slot == -1 is padding and must not be written (RUNNER:76-81)
  * block 0 is vLLM's null block and is never a write target (PLAT:149-159)

Logical pool layout used by the oracle: ``[num_blocks, block_size, Hkv, D]``
(token-major).  The CUDA pool is ``[num_blocks, Hkv, block_size, D]``; tests
permute before comparing, contents must be bit-identical.
"""
from __future__ import annotations

import torch


def kv_write_ref(k, v, k_pool, v_pool, slot_mapping):
    """k, v: [T, Hkv, D]; pools: [num_blocks, bs, Hkv, D]; slot_mapping: [T] int64 (-1 = skip)."""
    nb, bs, hkv, d = k_pool.shape
    kf = k_pool.view(nb * bs, hkv, d)
    vf = v_pool.view(nb * bs, hkv, d)
    slots = slot_mapping.reshape(-1).tolist()
    for t, s in enumerate(slots):
        if s < 0:
            continue
        kf[s] = k[t]
        vf[s] = v[t]


def rope_table(head_dim: int, max_pos: int, theta: float = 10000.0, scaling: dict | None = None):
    """fp32 cos/sin table [max_pos, head_dim/2]; 'llama3' scaling as published by Meta/HF."""
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float64) / head_dim))
    if scaling and scaling.get("rope_type", scaling.get("type")) == "llama3":
        factor = scaling["factor"]
        lo = scaling["low_freq_factor"]
        hi = scaling["high_freq_factor"]
        old = scaling["original_max_position_embeddings"]
        wavelen = 2 * torch.pi / inv_freq
        smooth = (old / wavelen - lo) / (hi - lo)
        scaled = torch.where(wavelen > old / lo, inv_freq / factor, inv_freq)
        mid = (1 - smooth) * inv_freq / factor + smooth * inv_freq
        is_mid = (wavelen <= old / lo) & (wavelen >= old / hi)
        inv_freq = torch.where(is_mid, mid, scaled)
    pos = torch.arange(max_pos, dtype=torch.float64)
    ang = torch.outer(pos, inv_freq)
    return ang.cos().float().contiguous(), ang.sin().float().contiguous()


def rope_ref(x, positions, cos, sin):
    """NeoX/HF-Llama rotate-half RoPE.  x: [T, H, D] bf16; positions [T]; tables fp32 -> bf16 out."""
    d = x.shape[-1]
    xf = x.float()
    c = cos[positions].unsqueeze(1)  # [T,1,D/2]
    s = sin[positions].unsqueeze(1)
    x1, x2 = xf[..., : d // 2], xf[..., d // 2 :]
    o1 = x1 * c - x2 * s
    o2 = x2 * c + x1 * s
    return torch.cat([o1, o2], dim=-1).to(x.dtype)


def rmsnorm_ref(x, w, eps):
    """x [T, H] bf16, w [H] bf16 -> bf16 (all fp32 math, one rounding at the end)."""
    xf = x.float()
    var = xf.pow(2).mean(-1, keepdim=True)
    y = xf * torch.rsqrt(var + eps)
    return (y * w.float()).to(x.dtype)


def silu_mul_ref(gate_up):
    """gate_up [T, 2I] bf16 (gate columns first) -> bf16 [T, I]."""
    i = gate_up.shape[-1] // 2
    g = gate_up[..., :i].float()
    u = gate_up[..., i:].float()
    return (g * torch.sigmoid(g) * u).to(gate_up.dtype)


def argmax_ref(logits):
    """First index of the maximum per row (ties -> lowest index)."""
    m = logits.max(dim=-1, keepdim=True).values
    idx = torch.arange(logits.shape[-1]).expand_as(logits)
    big = torch.full_like(idx, logits.shape[-1])
    return torch.where(logits == m, idx, big).min(dim=-1).values
