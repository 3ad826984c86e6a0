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


SHAPES = [
    (64, 6144, 4096, 0),     # cfg3 qkv
    (64, 4096, 4096, 0),     # o_proj (auto split-K)
    (64, 4096, 14336, 0),    # down_proj (auto split-K, long K)
    (64, 1024, 512, 1),      # no split
    (64, 1024, 512, 4),      # explicit split, 1 k-block each (fp8)
    (4, 2048, 2048, 0),      # TinyLlama decode batch
    (1, 256, 128, 0),        # single row, single k-block
    (128, 640, 1152, 0),     # Qwen TP=8 o-proj-like, M=128
    (100, 16032, 256, 0),    # N not a multiple of 128 (vocab shard), ragged M
    (300, 384, 640, 0),      # M > 256: two m-tiles
]


@pytest.mark.parametrize("m,n,k,split", SHAPES)
@pytest.mark.parametrize("kind", ["fp8", "int8"])
def test_w8a8_gemm(ops, m, n, k, split, kind):
    g = torch.Generator().manual_seed(m * 7 + n + k)
    x = torch.randn(m, k, generator=g).bfloat16()
    w = (torch.randn(n, k, generator=g) * 0.02).bfloat16()
    xq, xs = quant_ref.quantize_act_per_token(x, kind)
    wq, ws = quant_ref.quantize_weight_per_channel(w, kind)
    bias = torch.randn(n, generator=g).bfloat16() if n % 256 == 0 else None
    res = torch.randn(m, n, generator=g).bfloat16() if k % 1024 == 0 else None
    want = quant_ref.qgemm_ref(xq, xs, wq, ws, bias, res)
    got = ops.gemm(xq.cuda(), wq.cuda(), xs.cuda(), ws.cuda(), None if bias is None else bias.cuda(),
                   None if res is None else res.cuda(), split_k=split)
    torch.cuda.synchronize()
    _assert_close(got, want, exact_int=(kind == "int8"))


@pytest.mark.parametrize("m,n,k,split", SHAPES[:4] + SHAPES[5:8])
def test_bf16_gemm(ops, m, n, k, split):
    g = torch.Generator().manual_seed(m + n + k)
    x = torch.randn(m, k, generator=g).bfloat16()
    w = (torch.randn(n, k, generator=g) * 0.02).bfloat16()
    want = quant_ref.gemm_ref(x, w)
    got = ops.gemm(x.cuda(), w.cuda(), split_k=split)
    torch.cuda.synchronize()
    _assert_close(got, want)


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


def test_gemm_split_k_is_deterministic(ops):
    g = torch.Generator().manual_seed(2)
    x = torch.randn(64, 4096, generator=g).bfloat16().cuda()
    w = torch.randn(4096, 4096, generator=g).bfloat16().cuda()
    a = ops.gemm(x, w, split_k=4)
    b = ops.gemm(x, w, split_k=4)
    assert torch.equal(a, b)
