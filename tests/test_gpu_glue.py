"""GPU parity: RMSNorm / SiLU*mul / per-token quantisation / embedding / argmax vs the oracle."""
import pytest
import torch

from oracle import kv_ref, quant_ref

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from vllm_neuron_b200 import ops as o

    o.device_check(0)
    return o


def _ulp_close(got, want, frac_ok=0.999):
    """bf16 results of fp32 math in a different summation order: identical except rare 1-ulp flips."""
    got, want = got.float().cpu(), want.float()
    err = (got - want).abs()
    assert (err <= want.abs() * 2.0**-7 + 1e-30).all()
    assert (got == want).float().mean().item() >= frac_ok


@pytest.mark.parametrize("t,h", [(64, 4096), (4, 2048), (3, 5120), (1, 8192), (7, 64)])
def test_rmsnorm(ops, t, h):
    g = torch.Generator().manual_seed(h)
    x = (torch.randn(t, h, generator=g) * 3).bfloat16()
    w = (1 + 0.1 * torch.randn(h, generator=g)).bfloat16()
    want = kv_ref.rmsnorm_ref(x, w, 1e-5)
    got = ops.rmsnorm(x.cuda(), w.cuda(), 1e-5)
    _ulp_close(got, want, 0.99)


@pytest.mark.parametrize("kind", ["fp8", "int8"])
def test_rmsnorm_quant(ops, kind):
    g = torch.Generator().manual_seed(4)
    x = torch.randn(64, 4096, generator=g).bfloat16()
    w = (1 + 0.1 * torch.randn(4096, generator=g)).bfloat16()
    q, s, y = ops.rmsnorm(x.cuda(), w.cuda(), 1e-5, quant=kind, want_bf16=True)
    # the quantiser must be bit-exact on the device's own bf16 activations
    qw, sw = quant_ref.quantize_act_per_token(y.cpu(), kind)
    assert torch.equal(s.cpu(), sw)
    assert torch.equal(q.cpu().view(torch.uint8), qw.view(torch.uint8))


@pytest.mark.parametrize("kind", ["fp8", "int8"])
def test_quantize_per_token_bit_exact(ops, kind):
    g = torch.Generator().manual_seed(8)
    x = (torch.randn(33, 1536, generator=g) * 5).bfloat16()
    x[3] = 0  # all-zero row -> scale 1
    q, s = ops.quantize_per_token(x.cuda(), kind)
    qw, sw = quant_ref.quantize_act_per_token(x, kind)
    assert torch.equal(s.cpu(), sw)
    assert torch.equal(q.cpu().view(torch.uint8), qw.view(torch.uint8))


def test_silu_mul(ops):
    g = torch.Generator().manual_seed(6)
    gu = (torch.randn(64, 2 * 14336, generator=g) * 2).bfloat16()
    want = kv_ref.silu_mul_ref(gu)
    got = ops.silu_mul(gu.cuda())
    _ulp_close(got, want, 0.98)
    q, s = ops.silu_mul(gu.cuda(), quant="fp8")
    qw, sw = quant_ref.quantize_act_per_token(got.cpu(), "fp8")
    assert torch.equal(s.cpu(), sw)
    assert torch.equal(q.cpu().view(torch.uint8), qw.view(torch.uint8))


def test_embedding_bit_exact(ops):
    g = torch.Generator().manual_seed(2)
    table = torch.randn(32000, 2048, generator=g).bfloat16()
    tok = torch.randint(0, 32000, (67,), generator=g)
    got = ops.embedding(tok.cuda(), table.cuda())
    assert torch.equal(got.cpu().view(torch.int16), table[tok].view(torch.int16))


def test_argmax_first_max(ops):
    g = torch.Generator().manual_seed(3)
    logits = torch.randn(64, 128256, generator=g)
    logits[5, 77] = logits[5, 90000] = 50.0  # tie -> lowest index
    logits[6, 128255] = 60.0
    logits[7, 0] = 60.0
    val, idx = ops.argmax(logits.cuda(), want_val=True)
    want = kv_ref.argmax_ref(logits)
    assert torch.equal(idx.cpu(), want)
    assert idx[5].item() == 77 and idx[6].item() == 128255 and idx[7].item() == 0
    assert torch.equal(val.cpu(), logits.max(dim=-1).values)
    assert torch.equal(ops.argmax(logits.cuda(), idx_offset=1000).cpu(), want + 1000)
