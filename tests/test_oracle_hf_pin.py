"""Pin the floating-point oracle (oracle/llama_ref.py and through it attn_ref, kv_ref, quant_ref) to HF transformers.

tests/golden/hf_golden.json holds logits and 64 greedy tokens of HF's own LlamaForCausalLM / Qwen2ForCausalLM (fp32, CPU,
eager attention, DynamicCache) for seeded random-init checkpoints -- the reference's accuracy method
(/root/reference/test/e2e/online/online_server_runner.py:95-146) as committed vectors; tests/golden/gen_from_hf.py wrote it.

Three gates per case:
 1. ALGORITHM: the oracle with fp32 activations (no bf16 rounding points) replays HF's 64 steps teacher-forced through its
    block-KV prefill path (random non-contiguous block ids) and must reproduce HF's top-8 logits to 2e-4 of the logit
    spread and every greedy token; the first decode steps are replayed again through the paged DECODE path.
 2. QUANTISED MODES: same against HF + per-token fake-quant hooks + per-channel de-quantised weights.
 3. ROUNDING NOISE: the oracle as the GPU tests use it (bf16 activations) against the same HF vectors -- the distance is
    bf16 rounding noise; it is measured, bounded, and tokens must agree wherever HF's own top-2 margin exceeds it.
"""
import dataclasses
import json
import os

import pytest
import torch

from oracle import index_ref, llama_ref
from vllm_neuron_b200.config import PRESETS
from vllm_neuron_b200.model import random_init_full_weights

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "hf_golden.json")
BS = 32
rope_scaling_override = {}  # {"value": None} makes _oracle drop a case's rope scaling (sensitivity check below)


def _golden():
    with open(GOLDEN) as f:
        return json.load(f)


def _oracle(case, mode, act_dtype, total_len):
    shape = PRESETS[case["preset"]]
    if case.get("rope_scaling"):
        shape = dataclasses.replace(shape, rope_scaling=rope_scaling_override.get("value", case["rope_scaling"]))
    w = random_init_full_weights(shape, seed=case["seed"], std=case["std"], resid_std=case["resid_std"])
    nblk = -(-total_len // BS)
    mml = nblk * BS
    nb = 2 * nblk + 3
    cfg = llama_ref.RefConfig(shape.hidden, shape.layers, shape.num_q_heads, shape.num_kv_heads, shape.head_dim,
                              shape.intermediate, shape.vocab, shape.rope_theta, shape.rope_scaling, shape.rms_eps,
                              shape.qkv_bias, "bf16" if mode == "plain" else mode, mml, BS, act_dtype)
    ref = llama_ref.RefLlama(cfg, w, nb)
    g = torch.Generator().manual_seed(99)
    blocks = (torch.randperm(nb - 1, generator=g)[:nblk] + 1).tolist()  # non-contiguous, never the null block
    return shape, ref, blocks, mml


def _replay_prefill(case, mode, act_dtype):
    """Teacher-forced replay of HF's 64 greedy steps as ONE block-KV prefill over prompt + tokens[:-1]; returns the
    logits rows that predicted each of the 64 tokens [64, V] (and the pieces for a decode-path continuation)."""
    rec = case["modes"][mode]
    prompt, toks = case["prompt"], rec["tokens"]
    seq = prompt + toks[:-1]
    shape, ref, blocks, mml = _oracle(case, mode, act_dtype, len(seq) + 1)
    row = index_ref.prefill_index(len(seq), blocks, 0, BS, mml)
    _, logits = ref.prefill(torch.tensor(seq), torch.tensor(row["block_table"]), torch.tensor(row["slot_mapping"]), 0,
                            all_logits=True)
    return logits[len(prompt) - 1:].float(), rec


def _compare(logits, rec, tol, what):
    """logits [n, V] vs the golden top-8 of the first n steps.  Returns (max |diff| / std, #tokens equal, #checked)."""
    n = logits.shape[0]
    worst, same, must = 0.0, 0, 0
    for i in range(n):
        ids = torch.tensor(rec["top_ids"][i])
        want = torch.tensor(rec["top_vals"][i])
        std = rec["logit_std"][i]
        err = ((logits[i][ids] - want).abs().max() / std).item()
        worst = max(worst, err)
        got = int(logits[i].argmax())
        same += got == rec["tokens"][i]
        if (want[0] - want[1]).item() > 4 * tol * std:  # outside the tolerance band the token is determined
            must += 1
            assert got == rec["tokens"][i], f"{what}: step {i} token {got} != HF {rec['tokens'][i]}"
    assert worst <= tol, f"{what}: top-8 logits differ from HF by {worst:.2e} of the logit spread (bar {tol:.0e})"
    return worst, same, must


@pytest.mark.parametrize("key", ["tiny-test", "tiny-qwen"])
def test_oracle_algorithm_matches_hf_fp32(key):
    case = _golden()[key]
    logits, rec = _replay_prefill(case, "plain", torch.float32)
    worst, same, must = _compare(logits, rec, 2e-4, f"{key} prefill path")
    print(f"[{key}] fp32 oracle vs HF {case['arch']}: max top-8 logit diff {worst:.2e} of spread, {same}/64 tokens equal")
    assert same == 64
    # full first-step logits row (these shapes store it): every vocabulary entry, not only the top-8
    full = torch.tensor(rec["logits0"])
    assert ((logits[0] - full).abs().max() / rec["logit_std"][0]).item() <= 2e-4


@pytest.mark.parametrize("key", ["tiny-test", "tiny-qwen"])
def test_oracle_decode_path_matches_hf_fp32(key):
    """prefill(prompt) then 12 steps through the paged DECODE attention (block-table gather per sequence, batch of 2: the
    same sequence twice at different block ids), teacher-forced with HF's tokens."""
    case = _golden()[key]
    rec = case["modes"]["plain"]
    prompt, toks = case["prompt"], rec["tokens"]
    n = 12
    shape, ref, blocks, mml = _oracle(case, "plain", torch.float32, len(prompt) + n + 1)
    nblk = len(blocks)
    other = [b for b in range(1, 2 * nblk + 3) if b not in blocks][:nblk]
    rows = []
    for bl in (blocks, other):
        row = index_ref.prefill_index(len(prompt), bl, 0, BS, mml)
        _, lg = ref.prefill(torch.tensor(prompt), torch.tensor(row["block_table"]), torch.tensor(row["slot_mapping"]), 0)
        rows.append(lg.float()[0])
    assert torch.equal(rows[0], rows[1])
    got = [rows[0]]
    for i in range(1, n):
        ds = [index_ref.decode_index(len(prompt), i, bl, BS, mml) for bl in (blocks, other)]
        _, lg = ref.decode(torch.tensor([toks[i - 1]] * 2), torch.tensor([x for d in ds for x in d["position_ids"]]),
                           torch.tensor([x for d in ds for x in d["slot_mapping"]]),
                           torch.tensor([d["block_table"] for d in ds]), torch.tensor([d["full_context_len"] for d in ds]))
        assert torch.equal(lg[0], lg[1])  # block placement must not change a bit
        got.append(lg[0].float())
    worst, same, _ = _compare(torch.stack(got), rec, 2e-4, f"{key} decode path")
    print(f"[{key}] fp32 oracle DECODE path vs HF: max diff {worst:.2e} of spread, {same}/{n} tokens equal")
    assert same == n


def _row_errors(key, mode):
    """Error of every row of the teacher-forced replay (prompt rows, then the 64 generated steps) against HF's top-8."""
    case = _golden()[key]
    rec = case["modes"][mode]
    prompt, toks = case["prompt"], rec["tokens"]
    seq = prompt + toks[:-1]
    shape, ref, blocks, mml = _oracle(case, mode, torch.float32, len(seq) + 1)
    row = index_ref.prefill_index(len(seq), blocks, 0, BS, mml)
    _, logits = ref.prefill(torch.tensor(seq), torch.tensor(row["block_table"]), torch.tensor(row["slot_mapping"]), 0,
                            all_logits=True)
    logits = logits.float()
    ids = rec["prompt_top_ids"][:-1] + rec["top_ids"]   # the last prompt row IS step 0
    vals = rec["prompt_top_vals"][:-1] + rec["top_vals"]
    stds = rec["prompt_logit_std"][:-1] + rec["logit_std"]
    return [((logits[i][torch.tensor(ids[i])] - torch.tensor(vals[i])).abs().max() / stds[i]).item()
            for i in range(len(seq))]


@pytest.mark.parametrize("key", ["tiny-test", "tiny-qwen"])
def test_oracle_weight_only_matches_hf_dequantised_weights(key):
    """INT8 weight-only = HF holding the per-channel de-quantised weights: no activation rounding, exact to fp32 noise."""
    errs = _row_errors(key, "int8_wo")
    print(f"[{key}/int8_wo] oracle vs HF with de-quantised weights: max {max(errs):.2e} of spread over {len(errs)} rows")
    assert max(errs) <= 2e-4


@pytest.mark.parametrize("key", ["tiny-test", "tiny-qwen"])
@pytest.mark.parametrize("mode", ["fp8", "int8"])
def test_oracle_w8a8_matches_hf_fake_quant(key, mode):
    """W8A8 (per-token activation scales): HF + a fake-quant hook on every nn.Linear input + de-quantised weights.
    Rows agree to fp32 noise (<= 2e-4 of the logit spread) until the first activation that sits within fp32 rounding
    distance of an 8-bit rounding boundary rounds the other way in HF's fp32 GEMM chain than after the oracle's fp64
    accumulation; that one step (0.8 % of the row maximum for int8, up to 6 % of the element for e4m3) enters the KV
    cache and legitimately moves every later row by the quantisation-noise level.  So: the rows computed before any flip
    must agree exactly (the rule is the same), and every row must stay inside the noise band (nothing systematic)."""
    errs = _row_errors(key, mode)
    exact = sum(e <= 2e-4 for e in errs)
    print(f"[{key}/{mode}] oracle vs HF+fake-quant: {exact}/{len(errs)} rows agree to fp32 noise (<= 2e-4 of the spread); "
          f"rows after an 8-bit rounding flip: max {max(errs):.3f}, median {sorted(errs)[len(errs) // 2]:.3f} of the spread")
    assert exact >= 4, errs[:12]
    assert max(errs) <= 0.25


@pytest.mark.parametrize("key,bar", [("tiny-test", 0.05), ("tiny-qwen", 0.05)])
def test_bf16_oracle_rounding_noise_vs_hf(key, bar):
    """The oracle exactly as the GPU tests use it (bf16 activations / KV): distance to HF fp32 = bf16 rounding noise."""
    case = _golden()[key]
    logits, rec = _replay_prefill(case, "plain", torch.bfloat16)
    worst, same, must = _compare(logits, rec, bar, f"{key} bf16")
    print(f"[{key}] bf16 oracle vs HF fp32: max top-8 diff {worst:.3f} of spread, {same}/64 tokens equal, {must} determined")


def test_oracle_tinyllama_shape_matches_hf_fp32():
    """BASELINE configs[0] shape (22 layers, GQA 8, D 64, V 32000): fp32 oracle vs HF on prompt + 64 teacher-forced steps."""
    case = _golden()["tinyllama-1.1b"]
    logits, rec = _replay_prefill(case, "plain", torch.float32)
    worst, same, must = _compare(logits, rec, 5e-4, "tinyllama-1.1b")
    print(f"[tinyllama-1.1b] fp32 oracle vs HF: max top-8 logit diff {worst:.2e} of spread, {same}/64 tokens equal")
    assert same == 64


def test_oracle_llama3_rope_scaling_matches_hf_fp32():
    """Llama-3.1 frequency scaling (rope_type "llama3"), which the headline 8B / 70B models run and the plain-RoPE cases above
    never touch: HF's own LlamaRotaryEmbedding supplies the frequencies (28 of 32 scaled at this shape: untouched, smooth and
    divided-by-8 branches all present), the oracle's kv_ref.rope_table must land on the same logits to fp32 noise, through
    the block-KV prefill path over positions 0..153.  And the pin is not vacuous: the same oracle WITHOUT the scaling is far
    outside the bar."""
    case = _golden()["tiny-llama3rope"]
    assert case["rope_scaling"]["rope_type"] == "llama3"
    logits, rec = _replay_prefill(case, "plain", torch.float32)
    worst, same, must = _compare(logits, rec, 2e-4, "tiny-llama3rope prefill path")
    print(f"[tiny-llama3rope] fp32 oracle vs HF: max top-8 logit diff {worst:.2e} of spread, {same}/64 tokens equal")
    assert same == 64
    full = torch.tensor(rec["logits0"])
    assert ((logits[0] - full).abs().max() / rec["logit_std"][0]).item() <= 2e-4
    rope_scaling_override["value"] = None
    try:
        plain, _ = _replay_prefill(case, "plain", torch.float32)
    finally:
        rope_scaling_override.clear()
    drift = max(((plain[i][torch.tensor(rec["top_ids"][i])] - torch.tensor(rec["top_vals"][i])).abs().max()
                 / rec["logit_std"][i]).item() for i in range(plain.shape[0]))
    print(f"[tiny-llama3rope] the same oracle with plain RoPE is {drift:.3f} of the spread away")
    assert drift > 50 * 2e-4
