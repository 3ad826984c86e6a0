"""GPU parity: paged attention (decode and prefix-cached prefill) vs the torch-CPU oracle.

Tolerance from BASELINE.json north_star: attention outputs within 2e-2 relative error in bf16.
We assert relative Frobenius error <= 2e-2 and max |err| <= 2e-2 * max|ref| (both far looser than observed)."""
import pytest
import torch

from oracle import attn_ref
from parity_util import logical_to_device_pool, make_block_tables, rel_err

pytestmark = pytest.mark.gpu
TOL = 2e-2


@pytest.fixture(scope="module")
def ops():
    from vllm_neuron_b200 import ops as o

    o.device_check(0)
    return o


def _pools(nb, bs, hkv, d, g):
    kp = torch.randn(nb, bs, hkv, d, generator=g).bfloat16()
    vp = torch.randn(nb, bs, hkv, d, generator=g).bfloat16()
    return kp, vp


def _check(got, want):
    assert torch.isfinite(got.float()).all()
    assert rel_err(got, want) <= TOL
    assert (got.float() - want.float()).abs().max().item() <= TOL * want.float().abs().max().item() + 1e-3


@pytest.mark.parametrize(
    "b,hq,hkv,d,ctxs,quota",
    [
        (4, 32, 8, 128, [1, 31, 32, 33], 0),          # block-boundary edge cases
        (8, 32, 8, 128, [2048, 100, 777, 1024, 5, 64, 1999, 300], 0),  # ragged, cfg2 heads
        (3, 32, 4, 64, [500, 2048, 17], 4),           # TinyLlama heads (G=8, D=64), 4-block minimum share
        (5, 5, 1, 128, [96, 640, 33, 1, 250], 2),     # Qwen2.5-32B @ TP=8 (G=5)
        (2, 8, 1, 128, [4096, 3000], 32),             # 70B @ TP=8 (G=8), long context, 32-block minimum share
        (6, 4, 4, 128, [128] * 6, 1),                 # MHA (G=1), 1-block shares (every block its own partial)
        (64, 32, 8, 128, [2048 + 7 * i for i in range(64)], 0),  # cfg2/cfg3 batch: shares cut (sequence, head) pairs mid-way
        (7, 8, 2, 64, [0, 33, 0, 0, 1000, 64, 0], 3),  # empty sequences inside and at both ends of the batch
    ],
)
def test_decode_matches_oracle(ops, b, hq, hkv, d, ctxs, quota):
    g = torch.Generator().manual_seed(13)
    bs = 32
    max_blocks = max(-(-c // bs) for c in ctxs)
    nb = sum(-(-c // bs) for c in ctxs) + 2
    kp, vp = _pools(nb, bs, hkv, d, g)
    q = torch.randn(b, hq, d, generator=g).bfloat16()
    ctx = torch.tensor(ctxs, dtype=torch.int32)
    bt = make_block_tables(b, ctxs, bs, nb, max_blocks, g)
    scale = d**-0.5
    want = attn_ref.paged_attention_decode_ref(q, kp, vp, bt, ctx, scale)
    got = ops.paged_attn_decode(
        q.cuda(), logical_to_device_pool(kp).cuda(), logical_to_device_pool(vp).cuda(), bt.cuda(), ctx.cuda(), scale,
        min_quota=quota,
    )
    torch.cuda.synchronize()
    _check(got.cpu(), want)


def test_decode_row_ids_and_strided_q(ops):
    """Persistent block-table rows addressed through row_ids; q taken as a strided view of a packed qkv row."""
    g = torch.Generator().manual_seed(3)
    b, hq, hkv, d, bs = 4, 32, 8, 128, 32
    ctxs = [300, 40, 1000, 64]
    nb = 64
    kp, vp = _pools(nb, bs, hkv, d, g)
    bt_small = make_block_tables(b, ctxs, bs, nb, 32, g)
    rows = torch.tensor([5, 0, 3, 6], dtype=torch.int32)
    bt = torch.zeros((8, 32), dtype=torch.int32)
    bt[rows.long()] = bt_small
    qkv = torch.randn(b, (hq + 2 * hkv) * d, generator=g).bfloat16()
    q = qkv[:, : hq * d].reshape(b, hq, d)
    ctx = torch.tensor(ctxs, dtype=torch.int32)
    want = attn_ref.paged_attention_decode_ref(q, kp, vp, bt_small, ctx, 0.1)
    qd = qkv.cuda()[:, : hq * d].unflatten(1, (hq, d))
    got = ops.paged_attn_decode(qd, logical_to_device_pool(kp).cuda(), logical_to_device_pool(vp).cuda(), bt.cuda(),
                                ctx.cuda(), 0.1, row_ids=rows.cuda())
    _check(got.cpu(), want)


@pytest.mark.parametrize("kind", ["fp8", "int8"])
def test_decode_fused_per_token_quantisation_is_bit_exact(ops, kind):
    """The merge kernel's quantised output == the oracle's per-token quantiser applied to the kernel's own bf16 output."""
    from oracle import quant_ref

    g = torch.Generator().manual_seed(29)
    b, hq, hkv, d, bs = 9, 32, 8, 128, 32
    ctxs = [700, 1, 64, 2048, 333, 0, 95, 1500, 32]
    nb = sum(-(-c // bs) for c in ctxs) + 2
    kp, vp = _pools(nb, bs, hkv, d, g)
    bt = make_block_tables(b, ctxs, bs, nb, 64, g).cuda()
    q = torch.randn(b, hq, d, generator=g).bfloat16().cuda()
    ctx = torch.tensor(ctxs, dtype=torch.int32).cuda()
    kd, vd = logical_to_device_pool(kp).cuda(), logical_to_device_pool(vp).cuda()
    y = ops.paged_attn_decode(q, kd, vd, bt, ctx, 0.09)
    yq, ys = ops.paged_attn_decode(q, kd, vd, bt, ctx, 0.09, quant=kind)
    qw, sw = quant_ref.quantize_act_per_token(y.cpu().reshape(b, hq * d), kind)
    assert torch.equal(ys.cpu(), sw)
    assert torch.equal(yq.cpu().view(torch.uint8), qw.view(torch.uint8))


def test_decode_is_deterministic_and_pdl_invariant(ops):
    """Same bits with programmatic dependent launch on and off (the partition does not depend on timing)."""
    g = torch.Generator().manual_seed(31)
    b, hq, hkv, d, bs = 16, 32, 8, 128, 32
    ctxs = [100 * (i + 1) for i in range(b)]
    nb = sum(-(-c // bs) for c in ctxs) + 2
    kp, vp = _pools(nb, bs, hkv, d, g)
    bt = make_block_tables(b, ctxs, bs, nb, 50, g).cuda()
    q = torch.randn(b, hq, d, generator=g).bfloat16().cuda()
    ctx = torch.tensor(ctxs, dtype=torch.int32).cuda()
    kd, vd = logical_to_device_pool(kp).cuda(), logical_to_device_pool(vp).cuda()
    a = ops.paged_attn_decode(q, kd, vd, bt, ctx, 0.09)
    old = ops.set_pdl(False)
    try:
        c = ops.paged_attn_decode(q, kd, vd, bt, ctx, 0.09)
    finally:
        ops.set_pdl(old)
    assert torch.equal(a, c)


def test_decode_zero_context_row_is_zero(ops):
    g = torch.Generator().manual_seed(1)
    kp, vp = _pools(8, 32, 2, 128, g)
    q = torch.randn(2, 8, 128, generator=g).bfloat16()
    bt = torch.tensor([[1, 2], [0, 0]], dtype=torch.int32)
    ctx = torch.tensor([40, 0], dtype=torch.int32)
    got = ops.paged_attn_decode(q.cuda(), logical_to_device_pool(kp).cuda(), logical_to_device_pool(vp).cuda(), bt.cuda(),
                                ctx.cuda(), 0.1).cpu()
    assert (got[1] == 0).all()
    _check(got[:1], attn_ref.paged_attention_decode_ref(q[:1], kp, vp, bt[:1], ctx[:1], 0.1))


@pytest.mark.parametrize(
    "s_new,computed,hq,hkv,d",
    [(1, 0, 8, 2, 128), (33, 0, 8, 2, 128), (200, 96, 32, 8, 128), (64, 64, 32, 4, 64), (130, 1, 5, 1, 128), (1, 777, 8, 8, 64),
     # the tcgen05 kernel (>= 64 new tokens): several kv steps, odd GQA group (one CTA of the pair holds a single tile),
     # one query head per kv head, cached prefixes that are not multiples of the 128-token step, D = 64
     (513, 0, 8, 2, 128), (300, 1000, 10, 2, 128), (129, 37, 8, 8, 64), (128, 128, 4, 1, 128), (1024, 0, 32, 8, 128),
     (700, 333, 32, 4, 64)],
)
def test_prefill_matches_oracle(ops, s_new, computed, hq, hkv, d):
    g = torch.Generator().manual_seed(17)
    bs = 32
    total = s_new + computed
    nblk = -(-total // bs)
    nb = nblk + 3
    kp, vp = _pools(nb, bs, hkv, d, g)
    bt = make_block_tables(1, [total], bs, nb, nblk + 2, g)[0]
    q = torch.randn(s_new, hq, d, generator=g).bfloat16()
    scale = d**-0.5
    want = attn_ref.paged_attention_prefill_ref(q, kp, vp, bt, computed, scale)
    got = ops.paged_attn_prefill(q.cuda(), logical_to_device_pool(kp).cuda(), logical_to_device_pool(vp).cuda(), bt.cuda(),
                                 computed, scale)
    torch.cuda.synchronize()
    _check(got.cpu(), want)


def test_prefill_last_row_equals_decode(ops):
    """Size-independent property: the last query row of a prefill equals a decode step at the same context."""
    g = torch.Generator().manual_seed(23)
    hq, hkv, d, bs, total = 32, 8, 128, 32, 1500
    nblk = -(-total // bs)
    kp, vp = _pools(nblk + 2, bs, hkv, d, g)
    bt = make_block_tables(1, [total], bs, nblk + 2, nblk, g)
    q = torch.randn(100, hq, d, generator=g).bfloat16().cuda()
    kd, vd = logical_to_device_pool(kp).cuda(), logical_to_device_pool(vp).cuda()
    pre = ops.paged_attn_prefill(q, kd, vd, bt[0].cuda(), total - 100, 0.09)
    dec = ops.paged_attn_decode(q[-1:], kd, vd, bt.cuda(), torch.tensor([total], dtype=torch.int32, device="cuda"), 0.09)
    assert rel_err(pre[-1:].cpu(), dec.cpu()) < 1e-2


def test_prefill_prefix_consistency_full_size(ops):
    """Size-independent property at the headline head configuration (32 q heads, 8 kv heads, D = 128) and an 8k sequence,
    where the CPU oracle would take minutes: computing the last 4096 tokens against a 4096-token CACHED prefix gives the
    same rows as computing all 8192 tokens at once (different query tiles, kv-step counts and mask positions in the
    tcgen05 kernel), and the last row of either equals a decode step at context 8192."""
    g = torch.Generator().manual_seed(29)
    hq, hkv, d, bs, total, cached = 32, 8, 128, 32, 8192, 4096
    nblk = total // bs
    kp, vp = _pools(nblk + 2, bs, hkv, d, g)
    bt = make_block_tables(1, [total], bs, nblk + 2, nblk, g)
    q = torch.randn(total, hq, d, generator=g).bfloat16().cuda()
    kd, vd = logical_to_device_pool(kp).cuda(), logical_to_device_pool(vp).cuda()
    scale = d**-0.5
    full = ops.paged_attn_prefill(q, kd, vd, bt[0].cuda(), 0, scale)
    tail = ops.paged_attn_prefill(q[cached:].contiguous(), kd, vd, bt[0].cuda(), cached, scale)
    torch.cuda.synchronize()
    assert torch.isfinite(full.float()).all() and torch.isfinite(tail.float()).all()
    assert rel_err(tail.cpu(), full[cached:].cpu()) < 5e-3
    dec = ops.paged_attn_decode(q[-1:], kd, vd, bt.cuda(), torch.tensor([total], dtype=torch.int32, device="cuda"), scale)
    assert rel_err(full[-1:].cpu(), dec.cpu()) < 1e-2
    # causality: row r must not depend on keys > r -- perturb the last block's K/V, the first 8160 rows stay bit-identical
    kd2, vd2 = kd.clone(), vd.clone()
    last_blk = int(bt[0, nblk - 1])
    kd2[last_blk] += 1.0
    vd2[last_blk] -= 1.0
    full2 = ops.paged_attn_prefill(q, kd2, vd2, bt[0].cuda(), 0, scale)
    torch.cuda.synchronize()
    assert torch.equal(full2[: total - bs], full[: total - bs])
    assert not torch.equal(full2[total - bs:], full[total - bs:])
