"""Oracle (test infrastructure): one Llama/Qwen2 decoder forward over the block-KV cache.

Arithmetic PINNED to HF transformers (tests/golden/hf_golden.json, tests/test_oracle_hf_pin.py).  This is the CPU restatement of what the
reference's ``self.model(...)`` call (RUNNER:517-530 -> LOADER:336-348) does inside
NxDI: embedding -> L x (RMSNorm -> qkv -> RoPE -> KV write by slot_mapping -> paged
attention via block_tables -> o_proj -> +res -> RMSNorm -> gate/up -> SiLU*up -> down
-> +res) -> norm -> lm_head -> greedy argmax (on-device sampling with top_k=1,
RUNNER:1126-1128).  Inputs are exactly the tensors the reference's runner builds
(oracle/index_ref.py).  bf16 rounding points are placed where the CUDA path rounds
so greedy tokens can be compared step for step.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch

from . import attn_ref, kv_ref, quant_ref


@dataclass
class RefConfig:
    hidden: int
    layers: int
    num_q_heads: int
    num_kv_heads: int
    head_dim: int
    intermediate: int
    vocab: int
    rope_theta: float = 10000.0
    rope_scaling: dict | None = None
    rms_eps: float = 1e-5
    qkv_bias: bool = False
    quant: str = "bf16"  # "bf16" | "fp8" (W8A8 per-channel) | "int8" (W8A8) | "int8_wo"
    max_pos: int = 2048
    block_size: int = 32
    # dtype of activations and of the KV pools.  bf16 = the rounding points of the CUDA path (what the GPU tests compare
    # with); float32 = no intermediate rounding, used only by tests/test_oracle_hf_pin.py to pin the ALGORITHM (RoPE
    # convention, GQA mapping, norm placement, paged gather, causal mask, quantisation rule) against HF transformers.
    act_dtype: torch.dtype = torch.bfloat16
    quant_type: str = "per_channel_symmetric"  # or "per_tensor_symmetric" (LOADER:893-895)


class RefLinear:
    def __init__(self, w, quant: str, bias=None, tensor_rows=None):
        self.quant = quant
        self.bias = bias
        if quant == "bf16":
            self.w = w
        elif quant in ("fp8", "int8"):
            self.wq, self.ws = quant_ref.quantize_weight_per_channel(w, quant, tensor_rows)
        elif quant == "int8_wo":
            self.wq, self.ws = quant_ref.quantize_weight_per_channel(w, "int8", tensor_rows)
        else:
            raise ValueError(quant)

    def prepare_f64(self):
        """Widen the stored weight once (values unchanged) so repeated calls skip the per-call conversion."""
        if self.quant == "bf16":
            self.w = self.w.to(torch.float64)
        else:
            self.wq = self.wq.to(torch.float64)
        return self

    def __call__(self, x, residual=None, out_dtype=None):
        out_dtype = x.dtype if out_dtype is None else out_dtype
        if self.quant == "bf16":
            return quant_ref.gemm_ref(x, self.w, self.bias, residual, out_dtype)
        if self.quant == "int8_wo":
            return quant_ref.wo_gemm_ref(x, self.wq, self.ws, self.bias, residual, out_dtype)
        xq, xs = quant_ref.quantize_act_per_token(x, self.quant)
        return quant_ref.qgemm_ref(xq, xs, self.wq, self.ws, self.bias, residual, out_dtype)


class RefLlama:
    def __init__(self, cfg: RefConfig, weights: dict, num_blocks: int, quant_lm_head: bool = True):
        self.cfg = cfg
        q = cfg.quant
        self.embed = weights["embed"]
        self.final_norm = weights["final_norm"]
        pt = cfg.quant_type == "per_tensor_symmetric"
        assert pt or cfg.quant_type == "per_channel_symmetric", cfg.quant_type
        qd, kd, it = cfg.num_q_heads * cfg.head_dim, cfg.num_kv_heads * cfg.head_dim, cfg.intermediate
        tr_qkv = [(0, qd), (qd, qd + kd), (qd + kd, qd + 2 * kd)] if pt else None  # q / k / v are separate HF tensors
        tr_gu = [(0, it), (it, 2 * it)] if pt else None
        tr_h = [(0, cfg.hidden)] if pt else None
        self.lm_head = RefLinear(weights["lm_head"], q if quant_lm_head else "bf16",
                                 tensor_rows=[(0, cfg.vocab)] if pt else None)
        self.layers = []
        for lw in weights["layers"]:
            self.layers.append(
                {
                    "input_norm": lw["input_norm"],
                    "post_norm": lw["post_norm"],
                    "qkv": RefLinear(lw["wqkv"], q, lw.get("bqkv"), tr_qkv),
                    "o": RefLinear(lw["wo"], q, tensor_rows=tr_h),
                    "gate_up": RefLinear(lw["wgu"], q, tensor_rows=tr_gu),
                    "down": RefLinear(lw["wdown"], q, tensor_rows=tr_h),
                }
            )
        self.cos, self.sin = kv_ref.rope_table(cfg.head_dim, cfg.max_pos, cfg.rope_theta, cfg.rope_scaling)
        shp = (num_blocks, cfg.block_size, cfg.num_kv_heads, cfg.head_dim)
        self.k_pools = [torch.zeros(shp, dtype=cfg.act_dtype) for _ in range(cfg.layers)]
        self.v_pools = [torch.zeros(shp, dtype=cfg.act_dtype) for _ in range(cfg.layers)]

    def _forward(self, tokens, positions, slot_mapping, attn):
        c = self.cfg
        t = tokens.numel()
        x = self.embed[tokens.reshape(-1)].to(c.act_dtype)
        qd, kd = c.num_q_heads * c.head_dim, c.num_kv_heads * c.head_dim
        scale = c.head_dim**-0.5
        for li, L in enumerate(self.layers):
            h = kv_ref.rmsnorm_ref(x, L["input_norm"], c.rms_eps)
            qkv = L["qkv"](h)
            q = qkv[:, :qd].reshape(t, c.num_q_heads, c.head_dim)
            k = qkv[:, qd : qd + kd].reshape(t, c.num_kv_heads, c.head_dim)
            v = qkv[:, qd + kd :].reshape(t, c.num_kv_heads, c.head_dim)
            q = kv_ref.rope_ref(q, positions, self.cos, self.sin)
            k = kv_ref.rope_ref(k, positions, self.cos, self.sin)
            kv_ref.kv_write_ref(k, v, self.k_pools[li], self.v_pools[li], slot_mapping)
            a = attn(q, self.k_pools[li], self.v_pools[li], scale)
            x = L["o"](a.reshape(t, qd), residual=x)
            h2 = kv_ref.rmsnorm_ref(x, L["post_norm"], c.rms_eps)
            gu = L["gate_up"](h2)
            act = kv_ref.silu_mul_ref(gu)
            x = L["down"](act, residual=x)
        return x

    def _logits(self, x_rows):
        xf = kv_ref.rmsnorm_ref(x_rows, self.final_norm, self.cfg.rms_eps)
        return self.lm_head(xf, out_dtype=torch.float32)

    def decode(self, tokens, positions, slot_mapping, block_tables, ctx_lens):
        """tokens/positions/slot_mapping [B]; block_tables [B, max_blocks]; ctx_lens [B] = full_context_lens."""

        def attn(q, kp, vp, scale):
            return attn_ref.paged_attention_decode_ref(q, kp, vp, block_tables, ctx_lens, scale)

        x = self._forward(tokens, positions, slot_mapping, attn)
        logits = self._logits(x)
        return kv_ref.argmax_ref(logits), logits

    def prefill(self, prompt_tokens, block_table_row, slot_mapping_sliced, computed_len: int, all_logits: bool = False):
        """Batch-1 prefix-cached prefill (SCHED:116).  ``prompt_tokens`` is the FULL prompt as the
        reference passes it (RUNNER:721); only the uncached suffix is computed; ``slot_mapping_sliced``
        is the reference's slice that starts at the first uncached token (RUNNER:757-763).
        all_logits: return the logits of every computed position [S_new, V] (teacher-forced replay in the tests)
        instead of the last one (what the reference samples from, LOADER:350-365)."""
        s = prompt_tokens.numel()
        new = prompt_tokens[computed_len:]
        positions = torch.arange(computed_len, s)
        slots = slot_mapping_sliced[: s - computed_len]

        def attn(q, kp, vp, scale):
            return attn_ref.paged_attention_prefill_ref(q, kp, vp, block_table_row, computed_len, scale)

        x = self._forward(new, positions, slots, attn)
        logits = self._logits(x if all_logits else x[-1:])
        return kv_ref.argmax_ref(logits), logits
