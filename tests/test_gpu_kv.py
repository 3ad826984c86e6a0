"""GPU parity: device index construction and KV-cache writes are BIT-EXACT against the oracle
(which is pinned to the reference's own outputs in test_oracle_index.py)."""
import numpy as np
import pytest
import torch

from oracle import index_ref, kv_ref
from parity_util import device_to_logical_pool

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from vllm_neuron_b200 import ops as o

    o.device_check(0)
    return o


def test_decode_meta_matches_golden(ops, golden_index):
    """b200_build_decode_meta reproduces the reference's decode tensors (model_runner.py:765-832)."""
    n = 0
    for c in golden_index["index"]:
        if c["kind"] != "decode" or c["tkg"] or c["spec"]:
            continue
        bs, mml = c["block_size"], c["max_model_len"]
        b = len(c["reqs"])
        mb = mml // bs
        bt = torch.zeros((b + 3, mb), dtype=torch.int32)
        rows = torch.randperm(b + 3, generator=torch.Generator().manual_seed(b))[:b].to(torch.int32)  # scattered persistent rows
        assert len(set(rows.tolist())) == b
        ntok = torch.zeros(b + 3, dtype=torch.int32)
        for i, q in enumerate(c["reqs"]):
            r = int(rows[i])
            bt[r, : len(q["block_ids"])] = torch.tensor(q["block_ids"], dtype=torch.int32)
            ntok[r] = q["prompt_len"] + len(q["output_token_ids"])
        d = "cuda"
        slot = torch.empty(b, dtype=torch.int64, device=d)
        pos = torch.empty(b, dtype=torch.int64, device=d)
        ctx = torch.empty(b, dtype=torch.int32, device=d)
        ops.build_decode_meta(bt.to(d), rows.to(d), ntok.to(d), slot, pos, ctx, bs)
        want = c["out"]
        assert slot.cpu().tolist() == [s[0] for s in want["slot_mapping"]]
        assert pos.cpu().tolist() == [p[0] for p in want["position_ids"]]
        assert ctx.cpu().tolist() == [f[0] for f in want["full_context_lens"]]
        assert (ctx.cpu() - 1).tolist() == [f[0] for f in want["computed_context_lens"]]
        n += 1
    assert n >= 5


def test_prefill_meta_matches_golden(ops, golden_index):
    """b200_build_prefill_meta reproduces slot_mapping/position_ids of model_runner.py:739-763, 853-885."""
    n = 0
    for c in golden_index["index"]:
        if c["kind"] != "prefill":
            continue
        bs, mml = c["block_size"], c["max_model_len"]
        plen, ncomp = len(c["prompt_token_ids"]), c["num_computed_tokens"]
        bt_row = torch.tensor(c["out"]["block_tables"][0], dtype=torch.int32, device="cuda")
        slot = torch.empty(mml, dtype=torch.int64, device="cuda")
        pos = torch.empty(mml, dtype=torch.int64, device="cuda")
        ops.build_prefill_meta(bt_row, plen, ncomp, slot, pos, bs)
        assert slot.cpu().tolist() == c["out"]["slot_mapping"][0]
        # reference positions are range(prompt_len) for the FULL prompt; ours are for the uncached suffix
        want_pos = list(range(ncomp, plen)) + [0] * (mml - (plen - ncomp))
        assert pos.cpu().tolist() == want_pos
        n += 1
    assert n >= 10


def test_block_table_scatter_and_advance(ops):
    bt = torch.zeros((4, 8), dtype=torch.int32, device="cuda")
    upd = torch.tensor([[0, 0, 5], [0, 1, 9], [3, 7, 2], [9, 0, 1], [1, 8, 1]], dtype=torch.int32, device="cuda")
    ops.block_table_scatter(bt, upd)
    want = torch.zeros((4, 8), dtype=torch.int32)
    want[0, 0], want[0, 1], want[3, 7] = 5, 9, 2  # out-of-range updates are dropped
    assert torch.equal(bt.cpu(), want)
    ntok = torch.tensor([3, 10, 7, 0], dtype=torch.int32, device="cuda")
    rows = torch.tensor([2, 0], dtype=torch.int32, device="cuda")
    toks = torch.zeros(2, dtype=torch.int64, device="cuda")
    ops.advance_decode(torch.tensor([11, 22], device="cuda"), toks, ntok, rows)
    assert toks.cpu().tolist() == [11, 22] and ntok.cpu().tolist() == [4, 10, 8, 0]


@pytest.mark.parametrize("hkv,d,t", [(8, 128, 64), (4, 64, 5), (1, 128, 300), (2, 64, 1)])
def test_kv_write_bit_exact(ops, hkv, d, t):
    g = torch.Generator().manual_seed(11)
    nb, bs = 40, 32
    k = torch.randn(t, hkv, d, generator=g).bfloat16()
    v = torch.randn(t, hkv, d, generator=g).bfloat16()
    slots = (torch.randperm((nb - 1) * bs, generator=g)[:t] + bs).to(torch.int64)  # never block 0
    slots[::7] = -1  # padding entries must be skipped
    kp = torch.randn(nb, bs, hkv, d, generator=g).bfloat16()
    vp = torch.randn(nb, bs, hkv, d, generator=g).bfloat16()
    kd = kp.permute(0, 2, 1, 3).contiguous().cuda()
    vd = vp.permute(0, 2, 1, 3).contiguous().cuda()
    kv_ref.kv_write_ref(k, v, kp, vp, slots)
    ops.kv_write(k.cuda(), v.cuda(), kd, vd, slots.cuda())
    assert torch.equal(device_to_logical_pool(kd).cpu().view(torch.int16), kp.view(torch.int16))
    assert torch.equal(device_to_logical_pool(vd).cpu().view(torch.int16), vp.view(torch.int16))
    assert torch.equal(kd[0].cpu().view(torch.int16), kp[0].permute(1, 0, 2).contiguous().view(torch.int16))  # null block untouched


@pytest.mark.parametrize("hq,hkv,d,t", [(32, 8, 128, 64), (32, 4, 64, 7), (5, 1, 128, 33), (8, 1, 128, 1)])
def test_rope_kv_write_bit_exact(ops, hq, hkv, d, t):
    g = torch.Generator().manual_seed(5)
    nb, bs, max_pos = 24, 32, 4096
    qkv = torch.randn(t, (hq + 2 * hkv) * d, generator=g).bfloat16()
    pos = torch.randint(0, max_pos, (t,), generator=g)
    slots = (torch.randperm((nb - 1) * bs, generator=g)[:t] + bs).to(torch.int64)
    if t > 3:
        slots[2] = -1
    scaling = {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
               "original_max_position_embeddings": 8192}
    cos, sin = kv_ref.rope_table(d, max_pos, 500000.0, scaling)
    kp = torch.zeros(nb, bs, hkv, d, dtype=torch.bfloat16)
    vp = torch.zeros(nb, bs, hkv, d, dtype=torch.bfloat16)
    q = qkv[:, : hq * d].reshape(t, hq, d)
    k = qkv[:, hq * d : (hq + hkv) * d].reshape(t, hkv, d)
    v = qkv[:, (hq + hkv) * d :].reshape(t, hkv, d)
    q_want = kv_ref.rope_ref(q, pos, cos, sin)
    kv_ref.kv_write_ref(kv_ref.rope_ref(k, pos, cos, sin), v, kp, vp, slots)
    kd = torch.zeros(nb, hkv, bs, d, dtype=torch.bfloat16, device="cuda")
    vd = torch.zeros_like(kd)
    qkv_d = qkv.cuda()
    ops.rope_kv_write(qkv_d, pos.cuda(), cos.cuda(), sin.cuda(), kd, vd, slots.cuda(), hq)
    q_got = qkv_d[:, : hq * d].reshape(t, hq, d).cpu()
    assert torch.equal(q_got.view(torch.int16), q_want.view(torch.int16))
    assert torch.equal(device_to_logical_pool(kd).cpu().view(torch.int16), kp.view(torch.int16))
    assert torch.equal(device_to_logical_pool(vd).cpu().view(torch.int16), vp.view(torch.int16))


def test_decode_meta_speculative_slots_match_golden(ops, golden_index):
    """num_spec = k: slot_mapping [B, k] = slots[0] + i, the reference's EAGLE / fused-speculation rows (model_runner.py:825-830)."""
    n = 0
    for c in golden_index["index"]:
        if c["kind"] != "decode" or c["tkg"] or not c["spec"]:
            continue
        bs, mml, k = c["block_size"], c["max_model_len"], c["spec"]
        b = len(c["reqs"])
        bt = torch.zeros((b + 1, mml // bs), dtype=torch.int32)
        ntok = torch.zeros(b + 1, dtype=torch.int32)
        for i, q in enumerate(c["reqs"]):
            bt[i, : len(q["block_ids"])] = torch.tensor(q["block_ids"], dtype=torch.int32)
            ntok[i] = q["prompt_len"] + len(q["output_token_ids"])
        rows = torch.arange(b + 1, dtype=torch.int32)  # last row: padding (0 tokens) -> -1 in every column
        d = "cuda"
        slot = torch.empty((b + 1, k), dtype=torch.int64, device=d)
        pos = torch.empty(b + 1, dtype=torch.int64, device=d)
        ctx = torch.empty(b + 1, dtype=torch.int32, device=d)
        ops.build_decode_meta(bt.to(d), rows.to(d), ntok.to(d), slot, pos, ctx, bs)
        assert slot.cpu().tolist() == c["out"]["slot_mapping"] + [[-1] * k]
        assert pos.cpu().tolist()[:b] == [p[0] for p in c["out"]["position_ids"]]
        want = [index_ref.decode_index(q["prompt_len"], len(q["output_token_ids"]), q["block_ids"], bs, mml,
                                       num_speculative_tokens=k)["slot_mapping"] for q in c["reqs"]]
        assert slot.cpu().tolist()[:b] == want
        n += 1
    assert n >= 5


def test_remask_spec_output_matches_golden(ops, golden_index):
    """b200_remask_spec_output == the reference's _remask_fused_spec_output (model_loader.py:308-334) on its own outputs."""
    assert len(golden_index["remask"]) >= 4
    for c in golden_index["remask"]:
        got = ops.remask_spec_output(torch.tensor(c["accepted"], dtype=torch.int64, device="cuda"),
                                     torch.tensor(c["next_pos"], dtype=torch.int64, device="cuda"),
                                     torch.tensor(c["positions"], dtype=torch.int64, device="cuda"))
        assert got.cpu().tolist() == c["masked"]
        assert got.cpu().tolist() == index_ref.remask_fused_spec_output(c["accepted"], c["next_pos"], c["positions"])
