"""Oracle (test infrastructure): block-KV paged attention, decode and prefix-cached prefill.

Arithmetic pinned through llama_ref to HF transformers' eager attention (tests/test_oracle_hf_pin.py; the reference
runs it inside NxDI).  The
*inputs* follow the reference's host contract exactly:
  block_tables [B, max_model_len // bs] right-padded with 0 (RUNNER:746-752, 812-817),
  full_context_lens = position + 1, computed_context_lens = position (RUNNER:786-787)
  for decode; computed_context_lens = num_computed_tokens for prefill (RUNNER:754).
Math (first principles): gather K,V through the block table, softmax(q K^T * scale
+ causal mask) V in fp32, GQA by repeating each kv head over Hq/Hkv query heads,
output rounded once to the input dtype.
"""
from __future__ import annotations

import torch


def gather_kv(pool, block_table_row, ctx_len: int):
    """pool [num_blocks, bs, Hkv, D] -> [ctx_len, Hkv, D] via the block table (bit-exact gather)."""
    nb, bs, hkv, d = pool.shape
    nblk = -(-ctx_len // bs)
    ids = block_table_row[:nblk].long()
    return pool[ids].reshape(nblk * bs, hkv, d)[:ctx_len]


def paged_attention_decode_ref(q, k_pool, v_pool, block_tables, ctx_lens, scale: float):
    """q [B, Hq, D]; ctx_lens [B] (= full_context_lens: includes the token just written) -> [B, Hq, D]."""
    b, hq, d = q.shape
    hkv = k_pool.shape[2]
    g = hq // hkv
    out = torch.empty_like(q)
    for i in range(b):
        n = int(ctx_lens[i])
        if n <= 0:
            out[i] = 0
            continue
        k = gather_kv(k_pool, block_tables[i], n).float()  # [n, Hkv, D]
        v = gather_kv(v_pool, block_tables[i], n).float()
        qi = q[i].float().view(hkv, g, d)
        s = torch.einsum("hgd,nhd->hgn", qi, k) * scale
        p = torch.softmax(s, dim=-1)
        o = torch.einsum("hgn,nhd->hgd", p, v)
        out[i] = o.reshape(hq, d).to(q.dtype)
    return out


def paged_attention_prefill_ref(q, k_pool, v_pool, block_table_row, computed_len: int, scale: float):
    """One sequence.  q [S_new, Hq, D] are the queries at absolute positions
    computed_len .. computed_len+S_new-1; K/V for ALL positions (cached prefix and the
    new tokens) are already in the pool.  Causal over the absolute positions."""
    s_new, hq, d = q.shape
    hkv = k_pool.shape[2]
    g = hq // hkv
    total = computed_len + s_new
    k = gather_kv(k_pool, block_table_row, total).float()
    v = gather_kv(v_pool, block_table_row, total).float()
    qf = q.float().view(s_new, hkv, g, d)
    s = torch.einsum("shgd,nhd->hgsn", qf, k) * scale
    qpos = torch.arange(s_new).view(-1, 1) + computed_len
    kpos = torch.arange(total).view(1, -1)
    s = s.masked_fill(kpos > qpos, float("-inf"))
    p = torch.softmax(s, dim=-1)
    o = torch.einsum("hgsn,nhd->shgd", p, v)
    return o.reshape(s_new, hq, d).to(q.dtype)
