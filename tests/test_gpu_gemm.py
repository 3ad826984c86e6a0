"""GPU parity: per-channel-scaled GEMMs (tcgen05) vs the oracle's quantize/dequantize GEMM.

Stated tolerances (north_star: "quantized GEMMs within stated per-channel FP8/INT8 tolerances"):
  * the quantised operands are produced by the ORACLE's quantiser, so both sides multiply the same codes;
  * INT8 W8A8: products and sums are exact integers -> result must match the oracle to 1 bf16 ulp
    (rel 2^-7 per element; only the final rounding/scale order can differ);
  * FP8 e4m3 W8A8 and bf16: fp32 accumulation in a different order -> |err| <= 1e-2 * rms(ref) + 2^-7 |ref|.
"""
import pytest
import torch

from oracle import quant_ref

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from vllm_neuron_b200 import ops as o

    o.device_check(0)
    return o


def _assert_close(got, want, exact_int=False):
    got, want = got.float().cpu(), want.float()
    assert torch.isfinite(got).all()
    rms = want.pow(2).mean().sqrt().item()
    tol = (0.0 if exact_int else 1e-2 * rms) + want.abs() * 2.0**-7 + 1e-6
    bad = (got - want).abs() > tol
    assert not bad.any(), f"{bad.sum().item()} / {bad.numel()} outside tolerance; max err {(got - want).abs().max().item()} rms {rms}"


# last column = split_k: cluster size along K for the decode schedule (M <= 128; 0 = auto), CTA cap for stream-K (M > 128)
SHAPES = [
    (64, 6144, 4096, 0),     # cfg3 qkv (auto: clusters of 4)
    (64, 4096, 4096, 0),     # o_proj (auto: clusters of 8)
    (64, 4096, 14336, 0),    # down_proj (long K)
    (64, 1024, 512, 1),      # no split: epilogue straight from TMEM
    (64, 1024, 512, 2),      # clusters of 2: 64-row slices
    (64, 1024, 512, 4),      # clusters of 4, one k-block per CTA (fp8)
    (33, 896, 2048, 8),      # clusters of 8; N, M ragged (tokens 33 / 8 per owner)
    (64, 896, 2048, 3),      # clusters of 3: 21/21/22 tokens per owner
    (5, 512, 2048, 7),       # fewer tokens than ranks: some owners finish nothing
    (4, 2048, 2048, 0),      # TinyLlama decode batch
    (1, 256, 128, 0),        # single row, single k-block
    (128, 640, 1152, 0),     # Qwen TP=8 o-proj-like, M=128
    (128, 640, 1152, 4),     # M=128 tile through the cluster reduction
    (100, 16032, 256, 0),    # N not a multiple of 128 (vocab shard), ragged M
    (300, 384, 640, 0),      # M > 128: persistent stream-K, two m-tiles
    (300, 384, 640, 5),      # stream-K over 5 CTAs: spans start and end mid-tile
    (200, 1024, 512, 32),    # stream-K, every tile has several contributors
]


@pytest.mark.parametrize("m,n,k,ctas", SHAPES)
@pytest.mark.parametrize("kind", ["fp8", "int8"])
def test_w8a8_gemm(ops, m, n, k, ctas, kind):
    g = torch.Generator().manual_seed(m * 7 + n + k)
    x = torch.randn(m, k, generator=g).bfloat16()
    w = (torch.randn(n, k, generator=g) * 0.02).bfloat16()
    xq, xs = quant_ref.quantize_act_per_token(x, kind)
    wq, ws = quant_ref.quantize_weight_per_channel(w, kind)
    bias = torch.randn(n, generator=g).bfloat16() if n % 256 == 0 else None
    res = torch.randn(m, n, generator=g).bfloat16() if k % 1024 == 0 else None
    want = quant_ref.qgemm_ref(xq, xs, wq, ws, bias, res)
    got = ops.gemm(xq.cuda(), wq.cuda(), xs.cuda(), ws.cuda(), None if bias is None else bias.cuda(),
                   None if res is None else res.cuda(), split_k=ctas)
    torch.cuda.synchronize()
    _assert_close(got, want, exact_int=(kind == "int8"))


@pytest.mark.parametrize("m,n,k,ctas", SHAPES[:5] + SHAPES[6:10] + SHAPES[12:14])
def test_bf16_gemm(ops, m, n, k, ctas):
    g = torch.Generator().manual_seed(m + n + k)
    x = torch.randn(m, k, generator=g).bfloat16()
    w = (torch.randn(n, k, generator=g) * 0.02).bfloat16()
    want = quant_ref.gemm_ref(x, w)
    got = ops.gemm(x.cuda(), w.cuda(), split_k=ctas)
    torch.cuda.synchronize()
    _assert_close(got, want)


@pytest.mark.parametrize("m,n,k,ctas", SHAPES[:10] + SHAPES[11:])
def test_int8_weight_only_gemm(ops, m, n, k, ctas):
    """bf16 activations x int8 per-channel weights (cfg4, Qwen2.5-32B INT8 weight-only): the int8 rows are widened to
    bf16 inside the kernel (exact), so the result is the bf16 GEMM on the dequantised codes -> same bar as bf16."""
    g = torch.Generator().manual_seed(m * 3 + n + k)
    x = torch.randn(m, k, generator=g).bfloat16()
    w = (torch.randn(n, k, generator=g) * 0.02).bfloat16()
    wq, ws = quant_ref.quantize_weight_per_channel(w, "int8")
    bias = torch.randn(n, generator=g).bfloat16() if n % 256 == 0 else None
    res = torch.randn(m, n, generator=g).bfloat16() if k % 1024 == 0 else None
    want = quant_ref.wo_gemm_ref(x, wq, ws, bias, res)
    got = ops.gemm(x.cuda(), wq.cuda(), None, ws.cuda(), None if bias is None else bias.cuda(),
                   None if res is None else res.cuda(), split_k=ctas)
    torch.cuda.synchronize()
    _assert_close(got, want)


def test_int8_weight_only_exact_on_integer_activations(ops):
    """Size-independent property: with small-integer bf16 activations every product and partial sum is an exact integer
    below 2^24, so the weight-only GEMM must reproduce the integer GEMM bit for bit (decode and stream-K schedules)."""
    g = torch.Generator().manual_seed(5)
    n, k = 1280, 1152
    w = torch.randint(-127, 128, (n, k), generator=g, dtype=torch.int8)
    for m in (64, 128, 300):
        a = torch.randint(-8, 9, (m, k), generator=g).to(torch.bfloat16)
        got = ops.gemm(a.cuda(), w.cuda(), out_dtype=torch.float32)
        want = (a.double() @ w.double().t()).float()
        assert torch.equal(got.cpu(), want), m


def test_fp32_out_and_linearity(ops):
    """Size-independent property at a BASELINE-size shape (lm_head shard): GEMM(a1+a2) == GEMM(a1)+GEMM(a2) in fp32 out
    for int8 codes (exact integer arithmetic), and fp32 output equals the oracle."""
    g = torch.Generator().manual_seed(9)
    m, n, k = 64, 16032, 4096
    a1 = torch.randint(-50, 50, (m, k), generator=g, dtype=torch.int8)
    a2 = torch.randint(-50, 50, (m, k), generator=g, dtype=torch.int8)
    w = torch.randint(-127, 128, (n, k), generator=g, dtype=torch.int8)
    o1 = ops.gemm(a1.cuda(), w.cuda(), out_dtype=torch.float32)
    o2 = ops.gemm(a2.cuda(), w.cuda(), out_dtype=torch.float32)
    o12 = ops.gemm((a1 + a2).cuda(), w.cuda(), out_dtype=torch.float32)
    assert torch.equal(o1 + o2, o12)
    want = (a1[:8].double() @ w.double().t()).float()
    assert torch.equal(o1[:8].cpu(), want)


def test_gemm_is_deterministic_and_stream_k_leaves_workspace_clean(ops):
    """Partial sums are added in fixed order (cluster rank / stream-K contributor) -> same bits every launch, with PDL on
    or off; the stream-K counters at the head of a shared workspace are zero again after every call (that is what lets
    one workspace serve a whole step)."""
    g = torch.Generator().manual_seed(2)
    w = torch.randn(4096, 4096, generator=g).bfloat16().cuda()
    for m in (64, 200):
        x = torch.randn(m, 4096, generator=g).bfloat16().cuda()
        ws = ops.gemm_workspace(m, 4096, 4096, 0, x.device)
        assert (ws is None) == (m <= 128)
        a = ops.gemm(x, w, workspace=ws)
        b = ops.gemm(x, w, workspace=ws)
        old = ops.set_pdl(False)
        try:
            c = ops.gemm(x, w, workspace=ws)
        finally:
            ops.set_pdl(old)
        assert torch.equal(a, b) and torch.equal(a, c)
        if ws is not None:
            assert int(ws[:4096].view(torch.int32).abs().sum()) == 0


def test_gemm_rejects_bad_split(ops):
    from vllm_neuron_b200._cabi import B200Error

    x = torch.zeros(64, 256, dtype=torch.bfloat16, device="cuda")
    w = torch.zeros(256, 256, dtype=torch.bfloat16, device="cuda")
    for bad in (9, 16, 8, -1):  # above 8 / more CTAs than k-blocks (256 bf16 = 4 k-blocks) / negative
        with pytest.raises(B200Error):
            ops.gemm(x, w, split_k=bad)


@pytest.mark.parametrize("kind", ["bf16", "fp8", "int8", "int8_wo"])
@pytest.mark.parametrize("m,i,k,ctas", [(64, 14336, 4096, 0), (4, 5632, 2048, 0), (64, 5632, 2048, 2), (100, 96, 512, 4), (64, 5632, 2048, 6),
                                         (128, 3584, 1024, 1), (300, 1792, 512, 0), (300, 1792, 512, 3)])
def test_gemm_silu_mul_matches_unfused(ops, kind, m, i, k, ctas):
    """gate/up GEMM with SiLU*mul in the epilogue == b200_gemm (on the same interleaved rows) followed by b200_silu_mul,
    bit for bit (same tiles, same stream-K cuts, fixed fix-up order), and within the GEMM tolerance of the oracle.  The per-group
    maxima it emits are exact and give the per-token scale of the oracle's quantiser."""
    from oracle import kv_ref

    g = torch.Generator().manual_seed(m + i + k)
    x = torch.randn(m, k, generator=g).bfloat16()
    w = (torch.randn(2 * i, k, generator=g) * 0.03).bfloat16()
    if kind == "bf16":
        xq, xs, wq, ws = x, None, w, None
        want_gu = quant_ref.gemm_ref(x, w)
    elif kind == "int8_wo":
        wq, ws = quant_ref.quantize_weight_per_channel(w, "int8")
        xq, xs = x, None
        want_gu = quant_ref.wo_gemm_ref(x, wq, ws)
    else:
        xq, xs = quant_ref.quantize_act_per_token(x, kind)
        wq, ws = quant_ref.quantize_weight_per_channel(w, kind)
        want_gu = quant_ref.qgemm_ref(xq, xs, wq, ws)
    want = kv_ref.silu_mul_ref(want_gu)
    dev = lambda t: None if t is None else t.cuda()  # noqa: E731
    wi, si = ops.interleave_gate_up(dev(wq), dev(ws))
    # unfused twin on the SAME row order (same tiles, same stream-K cuts), de-interleaved before SiLU*mul
    gu_i = ops.gemm(dev(xq), wi, dev(xs), si, split_k=ctas)
    idx = torch.arange(i).view(-1, 16)
    perm = torch.cat([idx, idx + i], dim=1).reshape(-1).cuda()
    gu = torch.empty_like(gu_i)
    gu[:, perm] = gu_i
    unfused = ops.silu_mul(gu)
    act, group_amax = ops.gemm_silu_mul(dev(xq), wi, dev(xs), si, split_k=ctas)
    torch.cuda.synchronize()
    assert torch.equal(act, unfused)
    got, wantf = act.float().cpu(), want.float()
    rms = wantf.pow(2).mean().sqrt().item()
    assert ((got - wantf).abs() <= 2e-2 * rms + wantf.abs() * 2.0**-6 + 1e-6).all()
    assert group_amax.shape == (m, 2 * i // 32)
    assert torch.equal(group_amax.cpu(), act.float().abs().view(m, -1, 16).max(dim=2).values.cpu())
    for qk in ("fp8", "int8"):
        q1, s1 = ops.quantize_per_token(act, qk, group_amax=group_amax)
        qw, sw = quant_ref.quantize_act_per_token(act.cpu(), qk)
        assert torch.equal(s1.cpu(), sw)
        assert torch.equal(q1.cpu().view(torch.uint8), qw.view(torch.uint8))


@pytest.mark.parametrize("kind", ["bf16", "fp8", "int8_wo"])
@pytest.mark.parametrize("m,hq,hkv,d,k,split", [
    (64, 32, 8, 128, 4096, 0),   # cfg3 qkv: auto split (cluster of 5) -> fused epilogue
    (64, 4, 1, 128, 4096, 0),    # cfg3 TP=8 shard
    (37, 8, 2, 64, 256, 2),      # D=64: two heads per 128-row tile, q/k/v heads mixed inside one warp; ragged M
    (128, 10, 2, 64, 640, 4),    # tiny-qwen traits (GQA 5, bias) at M=128
    (64, 8, 2, 64, 256, 1),      # no split: library falls back to GEMM + rope kernel
    (200, 8, 2, 64, 256, 0),     # prefill-sized (stream-K): fallback
])
def test_gemm_rope_kv_write_matches_two_launches(ops, kind, m, hq, hkv, d, k, split):
    """b200_gemm_rope_kv_write == b200_gemm followed by b200_rope_kv_write, bit for bit: the qkv row (rotated q,
    un-rotated k, v) and both KV pools (slots < 0 skipped, untouched blocks stay zero)."""
    g = torch.Generator().manual_seed(5)
    n, bs, nb, max_pos = (hq + 2 * hkv) * d, 32, 40, 512
    w = torch.randn(n, k, generator=g) * 0.05
    a = torch.randn(m, k, generator=g)
    bias = (torch.randn(n, generator=g) * 0.1).bfloat16().cuda() if hq == 10 else None
    if kind == "bf16":
        aq, a_s, wq, w_s = a.bfloat16().cuda(), None, w.bfloat16().cuda(), None
    elif kind == "fp8":
        wq, w_s = quant_ref.quantize_weight_per_channel(w.bfloat16(), "fp8")
        aq, a_s = quant_ref.quantize_act_per_token(a.bfloat16(), "fp8")
        aq, a_s, wq, w_s = aq.cuda(), a_s.cuda(), wq.cuda(), w_s.cuda()
    else:
        wq, w_s = quant_ref.quantize_weight_per_channel(w.bfloat16(), "int8")
        aq, a_s, wq, w_s = a.bfloat16().cuda(), None, wq.cuda(), w_s.cuda()
    pos = torch.randint(0, max_pos, (m,), generator=g)
    slots = (torch.randperm((nb - 1) * bs, generator=g)[:m] + bs)
    slots[::7] = -1  # padding rows
    pos[3] = max_pos + 5  # clamped
    pos, slots = pos.cuda(), slots.cuda()
    inv = 1.0 / (10000.0 ** (torch.arange(0, d, 2).float() / d))
    ang = torch.arange(max_pos).float()[:, None] * inv[None, :]
    cos, sin = ang.cos().cuda().contiguous(), ang.sin().cuda().contiguous()

    def pools():
        return (torch.zeros((nb, hkv, bs, d), dtype=torch.bfloat16, device="cuda"),
                torch.zeros((nb, hkv, bs, d), dtype=torch.bfloat16, device="cuda"))

    kp1, vp1 = pools()
    ref = ops.gemm(aq, wq, a_s, w_s, bias, split_k=split)
    ops.rope_kv_write(ref, pos, cos, sin, kp1, vp1, slots, hq)
    kp2, vp2 = pools()
    got = ops.gemm_rope_kv_write(aq, wq, pos, cos, sin, kp2, vp2, slots, hq, a_s, w_s, bias, split_k=split)
    torch.cuda.synchronize()
    assert torch.equal(got.view(torch.int16), ref.view(torch.int16))
    assert torch.equal(kp2.view(torch.int16), kp1.view(torch.int16)) and torch.equal(vp2.view(torch.int16), vp1.view(torch.int16))
    assert kp1.abs().sum().item() > 0 and (kp2[0] == 0).all()


@pytest.mark.parametrize("kind", ["fp8", "int8"])
@pytest.mark.parametrize("m,i,k", [(64, 14336, 4096), (1, 512, 256), (33, 1792, 512), (128, 3584, 1024), (17, 7168, 2048)])
def test_gemm_silu_mul_quant_fused_equals_two_calls(ops, kind, m, i, k):
    """b200_gemm_silu_mul_quant (SiLU*mul + per-token quantisation inside the gate/up GEMM, CTAs meeting at a grid barrier)
    == b200_gemm_silu_mul + b200_quantize_per_token, bit for bit, launch after launch on the same state buffer."""
    g = torch.Generator().manual_seed(m * 7 + i)
    x = (torch.randn(m, k, generator=g) * 0.7).bfloat16()
    w = (torch.randn(2 * i, k, generator=g) * 0.05).bfloat16()
    xq, xs = quant_ref.quantize_act_per_token(x, kind)
    wq, ws = quant_ref.quantize_weight_per_channel(w, kind)
    wi, si = ops.interleave_gate_up(wq.cuda(), ws.cuda())
    state = torch.zeros(1040, dtype=torch.uint8, device="cuda")
    act, gam = ops.gemm_silu_mul(xq.cuda(), wi, xs.cuda(), si)
    want_q, want_s = ops.quantize_per_token(act, kind, group_amax=gam)
    for rep in range(3):  # the state carries over: the epoch-tagged maxima of earlier launches must never leak
        xin = xq.cuda() if rep != 1 else torch.zeros_like(xq).cuda()
        got = ops.gemm_silu_mul_quant(xin, wi, xs.cuda(), si, None, state, kind)
        if got is None:
            pytest.skip("shape does not run as one resident wave")
        torch.cuda.synchronize()
        if rep == 1:
            assert torch.all(got[1] == 1.0) and torch.all(got[0].view(torch.uint8).int() % 128 == 0)  # all-zero rows: scale 1
            continue
        assert torch.equal(got[1], want_s)
        assert torch.equal(got[0].view(torch.uint8), want_q.view(torch.uint8))
