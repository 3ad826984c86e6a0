"""GPU: b200_paged_attn_decode_static, the decode-attention entry point for callers that cannot allocate.

Who needs it.  b200_paged_attn_decode asks the caller for a workspace (the flash-decoding partials).  A host that cannot
call cudaMalloc has nothing to pass: a language binding that may only forward tensors its framework already owns, a
plugin loaded after the serving process gave all spare device memory to the KV pool, or a runtime that forbids
allocation on its hot path.  The static variant keeps the partials in a 32 MiB array inside the library image instead.

What is checked, through the raw C ABI (ctypes, no ops wrapper): on two calls in a row, with different inputs so that the
second call runs over the first call's leftovers, the static variant returns exactly the bits of the workspace variant
(same kernels, same schedule; the tolerance of the other attention tests, 2e-2, is therefore met with zero error), the
second call leaves the device's free memory untouched, and an oversized shape is refused with a message that names the
limit.  The file sorts last on purpose: a failure here must not hide other tests under `pytest -x`.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

BS = 32


@pytest.fixture(scope="module")
def lib():
    from vllm_neuron_b200 import _cabi

    lib = _cabi.load()
    assert lib.b200_device_check(0) == 0, lib.b200_last_error()
    torch.cuda.set_device(0)
    return lib


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _case(seed, b, hq, hkv, d, ctxs):
    """Device pools [nb][Hkv][32][D], a shuffled block table (block 0 = null block, never used), q and ctx_lens."""
    g = torch.Generator().manual_seed(seed)
    max_blocks = max(1, max(-(-c // BS) for c in ctxs))
    need = sum(-(-c // BS) for c in ctxs)
    nb = need + 3
    kp = torch.randn(nb, hkv, BS, d, generator=g).bfloat16().cuda()
    vp = torch.randn(nb, hkv, BS, d, generator=g).bfloat16().cuda()
    q = torch.randn(b, hq, d, generator=g).bfloat16().cuda()
    perm = (torch.randperm(nb - 1, generator=g) + 1).to(torch.int32)
    bt = torch.zeros(b, max_blocks, dtype=torch.int32)
    at = 0
    for i, c in enumerate(ctxs):
        n = -(-c // BS)
        bt[i, :n] = perm[at : at + n]
        at += n
    return dict(q=q, kp=kp, vp=vp, bt=bt.cuda(), ctx=torch.tensor(ctxs, dtype=torch.int32).cuda(), nb=nb, b=b, hq=hq,
                hkv=hkv, d=d, mb=max_blocks, scale=d**-0.5)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _run_workspace(lib, c, quota):
    n = lib.b200_paged_attn_decode_workspace_bytes(c["b"], c["hq"], c["hkv"], c["d"], c["mb"], quota)
    ws = torch.empty(n, dtype=torch.uint8, device="cuda")
    out = torch.full((c["b"], c["hq"], c["d"]), float("nan"), dtype=torch.bfloat16, device="cuda")
    rc = lib.b200_paged_attn_decode(
        _p(c["q"]), c["q"].stride(0), _p(c["kp"]), _p(c["vp"]), c["nb"], _p(c["bt"]), c["bt"].stride(0), None, _p(c["ctx"]),
        _p(out), None, None, 0, _p(ws), n, c["b"], c["hq"], c["hkv"], c["d"], BS, c["mb"], quota, c["scale"], _stream())
    assert rc == 0, lib.b200_last_error()
    torch.cuda.synchronize()
    return out


def _run_static(lib, c, quota):
    out = torch.full((c["b"], c["hq"], c["d"]), float("nan"), dtype=torch.bfloat16, device="cuda")
    rc = lib.b200_paged_attn_decode_static(
        _p(c["q"]), c["q"].stride(0), _p(c["kp"]), _p(c["vp"]), c["nb"], _p(c["bt"]), c["bt"].stride(0), None, _p(c["ctx"]),
        _p(out), None, None, 0, c["b"], c["hq"], c["hkv"], c["d"], BS, c["mb"], quota, c["scale"], _stream())
    assert rc == 0, lib.b200_last_error()
    torch.cuda.synchronize()
    return out


def test_static_scratch_matches_workspace_variant_twice_in_a_row(lib):
    assert lib.b200_paged_attn_decode_static_max_bytes() == 32 << 20
    # first call: long ragged contexts, many partials per (sequence, head); it also does the per-device address look-up
    first = _case(101, 8, 32, 8, 128, [2048, 100, 777, 1024, 5, 64, 1999, 300])
    # second call: other heads / head size, shorter contexts, empty sequences -> fewer partials than the leftovers
    second = _case(202, 7, 8, 2, 64, [0, 33, 0, 0, 1000, 64, 0])
    want1 = _run_workspace(lib, first, 0)
    want2 = _run_workspace(lib, second, 3)

    got1 = _run_static(lib, first, 0)
    free_before = torch.cuda.mem_get_info()[0]
    allocated_before = torch.cuda.memory_allocated()
    out2 = torch.full((second["b"], second["hq"], second["d"]), float("nan"), dtype=torch.bfloat16, device="cuda")
    torch.cuda.synchronize()
    free_mid = torch.cuda.mem_get_info()[0]
    c = second
    rc = lib.b200_paged_attn_decode_static(
        _p(c["q"]), c["q"].stride(0), _p(c["kp"]), _p(c["vp"]), c["nb"], _p(c["bt"]), c["bt"].stride(0), None, _p(c["ctx"]),
        _p(out2), None, None, 0, c["b"], c["hq"], c["hkv"], c["d"], BS, c["mb"], 3, c["scale"], _stream())
    assert rc == 0, lib.b200_last_error()
    torch.cuda.synchronize()
    # the call itself took no device memory (out2 was allocated before free_mid was read)
    assert torch.cuda.mem_get_info()[0] == free_mid
    assert free_mid <= free_before and torch.cuda.memory_allocated() >= allocated_before

    for got, want in ((got1, want1), (out2, want2)):
        assert torch.isfinite(got.float()).all()
        assert torch.equal(got, want)  # same kernels and schedule: zero error, inside the 2e-2 of test_gpu_attn.py
    # and once more with the first inputs: leftovers of the small call under a big one
    assert torch.equal(_run_static(lib, first, 0), want1)


def test_static_scratch_refuses_shapes_beyond_its_array(lib):
    b, hq, hkv, d, mb = 64, 32, 8, 128, 1024  # 32k context at the headline batch: 2.2 GB of partials at the default quota
    need = lib.b200_paged_attn_decode_workspace_bytes(b, hq, hkv, d, mb, 0)
    assert need > lib.b200_paged_attn_decode_static_max_bytes()
    dummy = torch.zeros(16, dtype=torch.int32, device="cuda")
    launches = lib.b200_launch_count()
    rc = lib.b200_paged_attn_decode_static(
        _p(dummy), 2, _p(dummy), _p(dummy), 4, _p(dummy), mb, None, _p(dummy), _p(dummy), None, None, 0, b, hq, hkv, d, BS, mb,
        0, 1.0, _stream())
    assert rc == -3  # B200_ERR_WORKSPACE
    msg = lib.b200_last_error().decode()
    assert str(need) in msg and str(32 << 20) in msg
    assert lib.b200_launch_count() == launches  # nothing was launched
