"""Generate tests/golden/hf_golden.json: logits and greedy tokens of HF transformers' own LlamaForCausalLM /
Qwen2ForCausalLM, fp32 on CPU, for seeded random-init checkpoints of the shapes the parity tests use.

This is the reference's own accuracy method -- greedy generation with HF transformers on CPU as the judge of the
device path (/root/reference/test/e2e/online/online_server_runner.py:95-146) -- turned into committed vectors, because
neither weights nor network exist on the GPU box.  tests/test_oracle_hf_pin.py asserts oracle/llama_ref.py (and through
it attn_ref, kv_ref, quant_ref) against these vectors; the GPU tests then compare the CUDA path with that oracle.

What HF supplies here: embedding, RMSNorm, q/k/v/o projections, rotate-half RoPE, GQA repeat, causal softmax attention
over a DynamicCache, SiLU-gated MLP, residuals, final norm, lm_head -- none of it written by this repo.  The quantised
modes wrap HF's nn.Linear inputs in a fake-quant hook (per-token symmetric, the rule of oracle/quant_ref.py restated with
torch's own float8 cast) and load per-channel de-quantised weights: HF still supplies the whole network; only the
quantisation rule itself has no external reference (the reference forwards config keys to NxDI, LOADER:885-898).

    python tests/golden/gen_from_hf.py [--only tiny-test,tiny-qwen,tinyllama-1.1b]

Runs in this (GPU-less) authoring container with transformers 5.5.0; takes a few minutes for the TinyLlama shape.
"""
from __future__ import annotations

import argparse
import dataclasses
import json
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

TOPK = 8
STEPS = 64

# (key, preset, seed, std, resid_std, prompt_len, modes)
CASES = [
    ("tiny-test", "tiny-test", 3, 0.05, None, 40, ("plain", "int8_wo", "fp8", "int8")),
    ("tiny-qwen", "tiny-qwen", 5, 0.05, None, 37, ("plain", "int8_wo", "fp8", "int8")),
    ("tinyllama-1.1b", "tinyllama-1.1b", 3, 0.02, None, 48, ("plain",)),
    # Llama-3.1's frequency scaling (rope_type "llama3", what the headline 8B / 70B models run): the tiny shape with the
    # 3.1 factors and an original context of 64, so that positions 0..153 cross all three branches of the rule (wavelength
    # < 16 untouched, > 64 divided by 8, smooth in between) -- with the real 8192 a short golden sequence would not see it
    ("tiny-llama3rope", "tiny-test", 7, 0.05, None, 90, ("plain",)),
]
ROPE_SCALING = {"tiny-llama3rope": {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                                    "original_max_position_embeddings": 64}}


def fake_quant_rows(x: torch.Tensor, kind: str) -> torch.Tensor:
    """Per-row symmetric quantise -> de-quantise in fp32 (scale = amax / qmax, fp8 e4m3fn RNE saturating at 448 or
    int8 round-half-even clamped at 127)."""
    amax = x.abs().amax(dim=-1, keepdim=True)
    qmax = 448.0 if kind == "fp8" else 127.0
    scale = torch.where(amax > 0, amax / torch.full_like(amax, qmax), torch.ones_like(amax))
    y = x / scale
    if kind == "fp8":
        q = y.clamp(-448.0, 448.0).to(torch.float8_e4m3fn).float()
    else:
        q = torch.round(y).clamp(-127, 127)
    return q * scale


def build_hf(shape, weights, mode: str):
    """HF model in fp32 holding `weights` (mode plain) or their per-channel quantise/de-quantise image, with the
    per-token activation fake-quant hook on every nn.Linear for the W8A8 modes."""
    from parity_util import hf_state_dict
    from transformers import LlamaConfig, LlamaForCausalLM, Qwen2Config, Qwen2ForCausalLM

    kw = dict(hidden_size=shape.hidden, num_hidden_layers=shape.layers, num_attention_heads=shape.num_q_heads,
              num_key_value_heads=shape.num_kv_heads, intermediate_size=shape.intermediate, vocab_size=shape.vocab,
              rms_norm_eps=shape.rms_eps, rope_theta=shape.rope_theta, max_position_embeddings=4096,
              tie_word_embeddings=False, hidden_act="silu", attn_implementation="eager")
    if shape.rope_scaling is not None:
        kw["rope_scaling"] = dict(shape.rope_scaling)
    if shape.qkv_bias:
        assert shape.head_dim * shape.num_q_heads == shape.hidden
        cfg, cls = Qwen2Config(**kw), Qwen2ForCausalLM
    else:
        cfg, cls = LlamaConfig(head_dim=shape.head_dim, attention_bias=False, **kw), LlamaForCausalLM
    model = cls(cfg).float().eval()  # built on the CPU directly: the rotary buffers are initialised by HF itself
    sd = {k: v.float() for k, v in hf_state_dict(weights, shape).items()}
    if mode != "plain":
        wkind = "fp8" if mode == "fp8" else "int8"
        for k in list(sd):
            if k.endswith("_proj.weight") or k == "lm_head.weight":
                sd[k] = fake_quant_rows(sd[k], wkind)  # per output channel = per row of W[N, K]
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected and not missing, (missing, unexpected)
    if mode in ("fp8", "int8"):
        for mod in model.modules():
            if isinstance(mod, torch.nn.Linear):
                mod.register_forward_pre_hook(lambda m, a, _k=mode: (fake_quant_rows(a[0], _k),))
    return model


@torch.no_grad()
def run_case(shape, weights, mode, prompt):
    model = build_hf(shape, weights, mode)
    ids = torch.tensor([prompt])
    out = model(ids, use_cache=True)
    cache = out.past_key_values
    logits = out.logits[0, -1].float()
    rec = {"tokens": [], "top_vals": [], "top_ids": [], "logit_mean": [], "logit_std": []}
    rec["logits0"] = logits.tolist() if shape.vocab <= 2048 else None
    # every prompt position as well (row i predicts token i+1): the quantised modes are compared row by row, because one
    # 8-bit rounding flip in an early row legitimately moves every later row (see tests/test_oracle_hf_pin.py)
    pl = out.logits[0].float()
    ptv, pti = pl.topk(TOPK, dim=-1)
    rec["prompt_top_vals"], rec["prompt_top_ids"] = ptv.tolist(), pti.tolist()
    rec["prompt_logit_std"] = pl.std(dim=-1).tolist()
    for _ in range(STEPS):
        tv, ti = logits.topk(TOPK)
        tok = int(ti[0])
        rec["tokens"].append(tok)
        rec["top_vals"].append(tv.tolist())
        rec["top_ids"].append(ti.tolist())
        rec["logit_mean"].append(logits.mean().item())
        rec["logit_std"].append(logits.std().item())
        out = model(torch.tensor([[tok]]), past_key_values=cache, use_cache=True)
        cache = out.past_key_values
        logits = out.logits[0, -1].float()
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default=None)
    ap.add_argument("--out", default=os.path.join(HERE, "hf_golden.json"))
    args = ap.parse_args()
    import transformers

    from vllm_neuron_b200.config import PRESETS
    from vllm_neuron_b200.model import random_init_full_weights

    torch.manual_seed(0)
    golden = {}
    if os.path.exists(args.out):
        golden = json.load(open(args.out))
    golden["_meta"] = {"transformers": transformers.__version__, "torch": torch.__version__, "dtype": "float32",
                       "steps": STEPS, "topk": TOPK, "generator": "tests/golden/gen_from_hf.py"}
    only = set(args.only.split(",")) if args.only else None
    for key, preset, seed, std, resid_std, plen, modes in CASES:
        if only and key not in only:
            continue
        shape = PRESETS[preset]
        if key in ROPE_SCALING:
            shape = dataclasses.replace(shape, rope_scaling=ROPE_SCALING[key])
        weights = random_init_full_weights(shape, seed=seed, std=std, resid_std=resid_std)
        g = torch.Generator().manual_seed(1000 + seed)
        prompt = torch.randint(3, shape.vocab, (plen,), generator=g).tolist()
        golden[key] = {"rope_scaling": shape.rope_scaling, "preset": preset, "seed": seed, "std": std, "resid_std": resid_std, "prompt": prompt,
                       "arch": "Qwen2ForCausalLM" if shape.qkv_bias else "LlamaForCausalLM", "modes": {}}
        for mode in modes:
            rec = run_case(shape, weights, mode, prompt)
            golden[key]["modes"][mode] = rec
            m = [v[0] - v[1] for v in rec["top_vals"]]
            print(f"{key} [{mode}]: first tokens {rec['tokens'][:8]}, min top-2 margin / logit std = "
                  f"{min(a / s for a, s in zip(m, rec['logit_std'])):.4f}", flush=True)
    with open(args.out, "w") as f:
        json.dump(golden, f)
    print("wrote", args.out, os.path.getsize(args.out), "bytes")


if __name__ == "__main__":
    main()
