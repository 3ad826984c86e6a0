"""Tensor-level wrappers over the C ABI: validate, take raw pointers and the current CUDA stream, call.

Torch is plumbing here (device memory + streams); every op below runs one of the hand-written
sm_100a kernels in libb200_paged.so.  Nothing in this file computes on the host.
"""
from __future__ import annotations

import torch

from . import _cabi
from ._cabi import BF16, F32, FP8_E4M3, INT8, check

_QDT = {None: BF16, "bf16": BF16, "fp8": FP8_E4M3, "int8": INT8}
_TORCH_Q = {FP8_E4M3: torch.float8_e4m3fn, INT8: torch.int8}
INT8_WO = 4  # GEMM ab_dtype only: bf16 activations x int8 weights
_DTN = {BF16: "bf16", FP8_E4M3: "fp8", INT8: "int8", INT8_WO: "int8wo"}


def _stream():
    return torch.cuda.current_stream().cuda_stream


# ------------------------------------------------------------------ per-launch device timing (bench.py roofline leg)
_PROF: list | None = None


class kernel_timer:
    """with ops.kernel_timer() as t: ...eager calls...; t.summary() -> {name: (count, mean_ms)}.
    CUDA events are recorded on the launching stream right around each C-ABI call (not usable under graph capture)."""

    def __enter__(self):
        global _PROF
        self.records = _PROF = []
        return self

    def __exit__(self, *exc):
        global _PROF
        _PROF = None
        return False

    def summary(self):
        torch.cuda.synchronize()
        acc: dict[str, list] = {}
        for name, e0, e1 in self.records:
            acc.setdefault(name, []).append(e0.elapsed_time(e1))
        return {k: (len(v), sum(v) / len(v)) for k, v in acc.items()}


class _k:
    __slots__ = ("name", "e0")

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if _PROF is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        if _PROF is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            _PROF.append((self.name, self.e0, e1))
        return False


def _ptr(t):
    return 0 if t is None else t.data_ptr()


def _cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _cabi.B200Error("b200 ops need CUDA tensors (no CPU fallback)")


def device_check(device: int = 0):
    check(_cabi.load().b200_device_check(device), "b200_device_check")


def set_pdl(enabled: bool) -> bool:
    """Programmatic dependent launch on/off for every kernel of the library; returns the previous setting."""
    return bool(_cabi.load().b200_set_pdl(1 if enabled else 0))


def launch_count() -> int:
    """Kernel launches issued through libb200_paged.so by this process so far."""
    return int(_cabi.load().b200_launch_count())


# ------------------------------------------------------------------ index construction
def build_decode_meta(block_table, row_ids, num_tokens, slot_mapping, positions, ctx_lens, block_size: int):
    """block_table int32 [rows, max_blocks]; row_ids int32 [B] or None; num_tokens int32 [rows];
    outputs: slot_mapping int64 [B] (or [B, k]: k consecutive slots per row for speculation, RUNNER:825-830),
    positions int64 [B], ctx_lens int32 [B]."""
    _cuda(block_table, num_tokens, slot_mapping)
    assert block_table.dtype == torch.int32 and num_tokens.dtype == torch.int32
    assert slot_mapping.dtype == torch.int64 and positions.dtype == torch.int64 and ctx_lens.dtype == torch.int32
    b = positions.numel()
    num_spec = 1 if slot_mapping.dim() == 1 else slot_mapping.shape[1]
    assert slot_mapping.is_contiguous() and slot_mapping.numel() == b * num_spec
    with _k("build_decode_meta"):
        check(
            _cabi.load().b200_build_decode_meta(
                _ptr(block_table), block_table.stride(0), _ptr(row_ids), _ptr(num_tokens), _ptr(slot_mapping),
                _ptr(positions), _ptr(ctx_lens), b, block_table.shape[1], block_size, num_spec, _stream(),
            ),
            "b200_build_decode_meta",
        )


def build_prefill_meta(block_table_row, prompt_len: int, num_computed: int, slot_mapping, positions, block_size: int):
    _cuda(block_table_row, slot_mapping, positions)
    assert block_table_row.dtype == torch.int32 and slot_mapping.dtype == torch.int64 and positions.dtype == torch.int64
    check(
        _cabi.load().b200_build_prefill_meta(
            _ptr(block_table_row), prompt_len, num_computed, _ptr(slot_mapping), _ptr(positions),
            slot_mapping.numel(), block_size, _stream(),
        ),
        "b200_build_prefill_meta",
    )


def remask_spec_output(accepted, next_pos, positions):
    """LOADER:308-334: accepted int64 [B, T] 0-padded -> -1 past clamp(next_pos - positions, 0, T) tokens per row."""
    _cuda(accepted, next_pos, positions)
    b, t = accepted.shape
    assert accepted.dtype == next_pos.dtype == positions.dtype == torch.int64 and accepted.is_contiguous()
    out = torch.empty_like(accepted)
    check(_cabi.load().b200_remask_spec_output(_ptr(accepted), _ptr(next_pos), _ptr(positions), _ptr(out), b, t, _stream()),
          "b200_remask_spec_output")
    return out


def block_table_scatter(block_table, updates):
    """updates int32 [n, 3] = (row, col, block_id) on the device."""
    _cuda(block_table, updates)
    assert block_table.dtype == torch.int32 and updates.dtype == torch.int32 and updates.is_contiguous()
    n = updates.shape[0]
    check(
        _cabi.load().b200_block_table_scatter(
            _ptr(block_table), block_table.stride(0), _ptr(updates), n, block_table.shape[0], block_table.shape[1], _stream()
        ),
        "b200_block_table_scatter",
    )


def advance_decode(sampled, tokens_io, num_tokens, row_ids):
    _cuda(sampled, tokens_io, num_tokens)
    check(
        _cabi.load().b200_advance_decode(
            _ptr(sampled), _ptr(tokens_io), _ptr(num_tokens), _ptr(row_ids), sampled.numel(), _stream()
        ),
        "b200_advance_decode",
    )


# ------------------------------------------------------------------ KV cache
def kv_write(k, v, k_pool, v_pool, slot_mapping):
    """k, v [T, Hkv, D] bf16 (last two dims contiguous); pools [num_blocks, Hkv, bs, D]."""
    _cuda(k, v, k_pool, v_pool, slot_mapping)
    t, hkv, d = k.shape
    nb, hkv2, bs, d2 = k_pool.shape
    assert (hkv, d) == (hkv2, d2) and k.stride(2) == 1 and k.stride(1) == d and v.stride(1) == d
    assert k_pool.is_contiguous() and v_pool.is_contiguous() and slot_mapping.dtype == torch.int64
    check(
        _cabi.load().b200_kv_write(
            _ptr(k), _ptr(v), k.stride(0), v.stride(0), _ptr(k_pool), _ptr(v_pool), _ptr(slot_mapping),
            t, hkv, d, bs, nb, _stream(),
        ),
        "b200_kv_write",
    )


def rope_kv_write(qkv, positions, cos, sin, k_pool, v_pool, slot_mapping, num_q_heads: int):
    """qkv [T, (Hq+2Hkv)*D] bf16 contiguous; q is rotated in place, roped k and v go to the pools."""
    _cuda(qkv, positions, cos, sin, k_pool, v_pool, slot_mapping)
    nb, hkv, bs, d = k_pool.shape
    t = qkv.shape[0]
    assert qkv.is_contiguous() and qkv.shape[1] == (num_q_heads + 2 * hkv) * d and qkv.dtype == torch.bfloat16
    assert cos.dtype == torch.float32 and cos.shape[1] == d // 2 and cos.is_contiguous() and sin.is_contiguous()
    assert positions.dtype == torch.int64 and slot_mapping.dtype == torch.int64
    with _k("rope_kv_write"):
        check(
            _cabi.load().b200_rope_kv_write(
                _ptr(qkv), _ptr(positions), _ptr(cos), _ptr(sin), cos.shape[0], _ptr(k_pool), _ptr(v_pool),
                _ptr(slot_mapping), t, num_q_heads, hkv, d, bs, nb, _stream(),
            ),
            "b200_rope_kv_write",
        )


# ------------------------------------------------------------------ paged attention
def paged_attn_decode_workspace_bytes(b, hq, hkv, d, max_blocks, min_quota=0) -> int:
    return int(_cabi.load().b200_paged_attn_decode_workspace_bytes(b, hq, hkv, d, max_blocks, min_quota))


def paged_attn_decode(q, k_pool, v_pool, block_table, ctx_lens, scale, out=None, workspace=None, row_ids=None,
                      max_blocks=None, min_quota=0, quant=None):
    """q [B, Hq, D] bf16 (token stride arbitrary); block_table int32 [rows, max_blocks]; ctx_lens int32 [B].
    quant None -> bf16 out [B, Hq, D]; "fp8"/"int8" -> (q [B, Hq*D], scale [B]) quantised per token in the merge kernel."""
    _cuda(q, k_pool, v_pool, block_table, ctx_lens)
    b, hq, d = q.shape
    nb, hkv, bs, d2 = k_pool.shape
    assert d == d2 and q.stride(2) == 1 and q.stride(1) == d and q.dtype == torch.bfloat16
    assert block_table.dtype == torch.int32 and ctx_lens.dtype == torch.int32
    mb = block_table.shape[1] if max_blocks is None else min(max_blocks, block_table.shape[1])
    qd = _QDT[quant]
    oq = osc = None
    if qd == BF16:
        if out is None:
            out = torch.empty((b, hq, d), dtype=q.dtype, device=q.device)
        assert out.is_contiguous()
    else:
        oq = torch.empty((b, hq * d), dtype=_TORCH_Q[qd], device=q.device)
        osc = torch.empty((b,), dtype=torch.float32, device=q.device)
    if workspace is None:
        workspace = torch.empty(paged_attn_decode_workspace_bytes(b, hq, hkv, d, mb, min_quota), dtype=torch.uint8, device=q.device)
    with _k("paged_attn_decode"):
        check(
            _cabi.load().b200_paged_attn_decode(
                _ptr(q), q.stride(0), _ptr(k_pool), _ptr(v_pool), nb, _ptr(block_table), block_table.stride(0),
                _ptr(row_ids), _ptr(ctx_lens), _ptr(out), _ptr(oq), _ptr(osc), qd, _ptr(workspace),
                workspace.numel() * workspace.element_size(), b, hq, hkv, d, bs, mb, min_quota, float(scale), _stream(),
            ),
            "b200_paged_attn_decode",
        )
    return out if qd == BF16 else (oq, osc)


def paged_attn_prefill(q, k_pool, v_pool, block_table_row, computed_len: int, scale, out=None):
    """q [S_new, Hq, D] bf16; one sequence; K/V of all computed_len + S_new tokens already in the pools."""
    _cuda(q, k_pool, v_pool, block_table_row)
    s_new, hq, d = q.shape
    nb, hkv, bs, _ = k_pool.shape
    assert q.stride(2) == 1 and q.stride(1) == d and block_table_row.dtype == torch.int32
    if out is None:
        out = torch.empty((s_new, hq, d), dtype=q.dtype, device=q.device)
    with _k("paged_attn_prefill"):
        check(
            _cabi.load().b200_paged_attn_prefill(
                _ptr(q), q.stride(0), _ptr(k_pool), _ptr(v_pool), nb, _ptr(block_table_row), _ptr(out), s_new,
                computed_len, hq, hkv, d, bs, float(scale), _stream(),
            ),
            "b200_paged_attn_prefill",
        )
    return out


# ------------------------------------------------------------------ GEMM
def _ab_dtype(t):
    if t.dtype == torch.bfloat16:
        return BF16
    if t.dtype == torch.float8_e4m3fn:
        return FP8_E4M3
    if t.dtype == torch.int8:
        return INT8
    raise _cabi.B200Error(f"unsupported GEMM operand dtype {t.dtype}")


def _gemm_dtype(a, w):
    """ab_dtype code of a GEMM: both operands alike, or bf16 activations x int8 weights (weight-only)."""
    if a.dtype == torch.bfloat16 and w.dtype == torch.int8:
        return INT8_WO
    if a.dtype != w.dtype:
        raise _cabi.B200Error(f"GEMM operand dtypes {a.dtype} x {w.dtype} not supported")
    return _ab_dtype(a)


def gemm_decode_plan(m, n, k, ab_dtype, split_k=0, silu_mul=False, has_residual=False) -> dict:
    """The decode schedule the library picks for this shape on the current device (for benches / DESIGN notes)."""
    import ctypes

    out = (ctypes.c_int * 6)()
    check(_cabi.load().b200_gemm_decode_plan(m, n, k, ab_dtype, split_k, int(silu_mul), int(has_residual), out),
          "b200_gemm_decode_plan")
    return dict(zip(("split", "ctas_per_sm", "w_stages", "a_stages", "cap_ctas", "smem"), list(out)))


def gemm_workspace_bytes(m, n, k, ab_dtype, split_k=0) -> int:
    return int(_cabi.load().b200_gemm_workspace_bytes(m, n, k, ab_dtype, split_k))


def gemm_workspace(m, n, k, ab_dtype, device, split_k=0):
    """Zero-filled workspace (stream-K fix-up counters at its head must start at zero; every call leaves them zeroed).
    The decode schedule (M <= 128) needs none: returns None."""
    nbytes = gemm_workspace_bytes(m, n, k, ab_dtype, split_k)
    return torch.zeros(nbytes, dtype=torch.uint8, device=device) if nbytes else None


def gemm(a, w, a_scale=None, w_scale=None, bias=None, residual=None, out=None, out_dtype=torch.bfloat16, split_k=0,
         workspace=None):
    """out[M,N] = (a[M,K] @ w[N,K]^T) * a_scale[m] * w_scale[n] + bias[n] + residual[m,n]."""
    _cuda(a, w, a_scale, w_scale, bias, residual)
    m, k = a.shape
    n, k2 = w.shape
    assert k == k2 and a.is_contiguous() and w.is_contiguous()
    dt = _gemm_dtype(a, w)
    if out is None:
        out = torch.empty((m, n), dtype=out_dtype, device=a.device)
    assert out.is_contiguous() and out.shape == (m, n)
    odt = F32 if out.dtype == torch.float32 else BF16
    if residual is not None:
        assert residual.dtype == torch.bfloat16 and residual.is_contiguous() and residual.shape == (m, n)
    if workspace is None:
        workspace = gemm_workspace(m, n, k, dt, a.device, split_k)
    with _k(f"gemm_{_DTN[dt]}[{m}x{n}x{k}]"):
        check(
            _cabi.load().b200_gemm(
                _ptr(a), _ptr(w), _ptr(out), _ptr(a_scale), _ptr(w_scale), _ptr(bias), _ptr(residual), m, n, k, dt, odt,
                split_k, _ptr(workspace), 0 if workspace is None else workspace.numel() * workspace.element_size(), _stream(),
            ),
            "b200_gemm",
        )
    return out


def gemm_rope_kv_write(a, w, positions, cos, sin, k_pool, v_pool, slot_mapping, num_q_heads: int, a_scale=None,
                       w_scale=None, bias=None, split_k=0, workspace=None):
    """qkv[M, (Hq+2Hkv)*D] = gemm(a, w, ...) with rope_kv_write fused into the GEMM epilogue when the decode schedule
    allows (one launch), else GEMM + rope_kv_write inside the library (two launches); identical bits.  Returns qkv."""
    _cuda(a, w, a_scale, w_scale, bias, positions, cos, sin, k_pool, v_pool, slot_mapping)
    m, k = a.shape
    n, k2 = w.shape
    nb, hkv, bs, d = k_pool.shape
    assert k == k2 and a.is_contiguous() and w.is_contiguous() and n == (num_q_heads + 2 * hkv) * d
    assert cos.dtype == torch.float32 and cos.shape[1] == d // 2 and cos.is_contiguous() and sin.is_contiguous()
    assert positions.dtype == torch.int64 and slot_mapping.dtype == torch.int64
    dt = _gemm_dtype(a, w)
    out = torch.empty((m, n), dtype=torch.bfloat16, device=a.device)
    if workspace is None:
        workspace = gemm_workspace(m, n, k, dt, a.device, split_k)
    with _k(f"gemm_rope_kv_{_DTN[dt]}[{m}x{n}x{k}]"):
        check(
            _cabi.load().b200_gemm_rope_kv_write(
                _ptr(a), _ptr(w), _ptr(out), _ptr(a_scale), _ptr(w_scale), _ptr(bias), m, n, k, dt, split_k,
                _ptr(workspace), 0 if workspace is None else workspace.numel() * workspace.element_size(),
                _ptr(positions), _ptr(cos), _ptr(sin), cos.shape[0], _ptr(k_pool), _ptr(v_pool), _ptr(slot_mapping),
                num_q_heads, hkv, d, bs, nb, _stream(),
            ),
            "b200_gemm_rope_kv_write",
        )
    return out


def interleave_gate_up(w_gu, *others):
    """Row permutation expected by gemm_silu_mul: [gate(I); up(I)] -> groups of 32 rows = 16 gate rows followed by the
    16 up rows of the same channels.  Applied (load time) to the weight and to every per-row companion (scale, bias)."""
    i = w_gu.shape[0] // 2
    assert i % 16 == 0, "intermediate size must be a multiple of 16"
    idx = torch.arange(i, device=w_gu.device).view(-1, 16)
    perm = torch.cat([idx, idx + i], dim=1).reshape(-1)
    res = [w_gu[perm].contiguous()] + [None if o is None else o[perm].contiguous() for o in others]
    return res[0] if not others else tuple(res)


def gemm_silu_mul(a, w_gu_interleaved, a_scale=None, w_scale=None, bias=None, split_k=0, workspace=None):
    """act[M, I] = bf16(silu(bf16(gate)) * bf16(up)) with gate/up = the two halves of a @ w_gu^T (rows interleaved by
    interleave_gate_up); also returns group_amax [M, 2I/32] for quantize_per_token."""
    _cuda(a, w_gu_interleaved, a_scale, w_scale, bias)
    m, k = a.shape
    n2, k2 = w_gu_interleaved.shape
    assert k == k2 and a.is_contiguous() and w_gu_interleaved.is_contiguous()
    dt = _gemm_dtype(a, w_gu_interleaved)
    act = torch.empty((m, n2 // 2), dtype=torch.bfloat16, device=a.device)
    assert n2 % 32 == 0
    group_amax = torch.empty((m, n2 // 32), dtype=torch.float32, device=a.device)
    if workspace is None:
        workspace = gemm_workspace(m, n2, k, dt, a.device, split_k)
    with _k(f"gemm_silu_mul_{_DTN[dt]}[{m}x{n2}x{k}]"):
        check(
            _cabi.load().b200_gemm_silu_mul(
                _ptr(a), _ptr(w_gu_interleaved), _ptr(act), _ptr(a_scale), _ptr(w_scale), _ptr(bias), _ptr(group_amax),
                m, n2, k, dt, split_k, _ptr(workspace), 0 if workspace is None else workspace.numel() * workspace.element_size(),
                _stream(),
            ),
            "b200_gemm_silu_mul",
        )
    return act, group_amax


_UNSUPPORTED = -4  # B200_ERR_UNSUPPORTED


def gemm_silu_mul_quant(a, w_gu_interleaved, a_scale, w_scale, bias, state, quant: str):
    """gate/up GEMM + SiLU*mul + per-token quantisation in ONE launch (b200_gemm_silu_mul_quant): returns (q [M, I], scale [M])
    or None when the shape does not run as one resident wave of one CTA per weight tile (the caller then uses gemm_silu_mul
    + quantize_per_token, which produce the same bits).  `state`: 1040-byte uint8 device tensor, zeroed once."""
    _cuda(a, w_gu_interleaved, a_scale, w_scale, bias, state)
    m, k = a.shape
    n2, k2 = w_gu_interleaved.shape
    assert k == k2 and a.is_contiguous() and w_gu_interleaved.is_contiguous() and n2 % 32 == 0 and state.numel() >= 1040
    if m > 128:
        return None
    dt = _gemm_dtype(a, w_gu_interleaved)
    qd = _QDT[quant]
    q = torch.empty((m, n2 // 2), dtype=_TORCH_Q[qd], device=a.device)
    sc = torch.empty((m,), dtype=torch.float32, device=a.device)
    with _k(f"gemm_silu_mul_quant_{_DTN[dt]}[{m}x{n2}x{k}]"):
        rc = _cabi.load().b200_gemm_silu_mul_quant(_ptr(a), _ptr(w_gu_interleaved), _ptr(q), _ptr(sc), _ptr(a_scale), _ptr(w_scale),
                                                   _ptr(bias), _ptr(state), m, n2, k, dt, _stream())
    if rc == _UNSUPPORTED:
        return None
    check(rc, "b200_gemm_silu_mul_quant")
    return q, sc


# ------------------------------------------------------------------ glue
def rmsnorm(x, w, eps, quant=None, out_bf16=None, want_bf16=None):
    """Returns bf16 y, or (q, scale) for quant in {"fp8","int8"} (plus y when want_bf16)."""
    _cuda(x, w)
    t, h = x.shape
    assert x.is_contiguous() and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
    qd = _QDT[quant]
    if qd == BF16:
        y = out_bf16 if out_bf16 is not None else torch.empty_like(x)
        with _k("rmsnorm"):
            check(_cabi.load().b200_rmsnorm(_ptr(x), _ptr(w), _ptr(y), 0, 0, t, h, float(eps), qd, _stream()), "b200_rmsnorm")
        return y
    q = torch.empty((t, h), dtype=_TORCH_Q[qd], device=x.device)
    s = torch.empty((t,), dtype=torch.float32, device=x.device)
    y = (out_bf16 if out_bf16 is not None else torch.empty_like(x)) if want_bf16 else None
    with _k("rmsnorm_quant"):
        check(_cabi.load().b200_rmsnorm(_ptr(x), _ptr(w), _ptr(y), _ptr(q), _ptr(s), t, h, float(eps), qd, _stream()), "b200_rmsnorm")
    return (q, s, y) if want_bf16 else (q, s)


def silu_mul(gate_up, quant=None):
    _cuda(gate_up)
    t, i2 = gate_up.shape
    i = i2 // 2
    assert gate_up.is_contiguous() and gate_up.dtype == torch.bfloat16
    qd = _QDT[quant]
    if qd == BF16:
        y = torch.empty((t, i), dtype=torch.bfloat16, device=gate_up.device)
        with _k("silu_mul"):
            check(_cabi.load().b200_silu_mul(_ptr(gate_up), _ptr(y), 0, 0, t, i, qd, _stream()), "b200_silu_mul")
        return y
    q = torch.empty((t, i), dtype=_TORCH_Q[qd], device=gate_up.device)
    s = torch.empty((t,), dtype=torch.float32, device=gate_up.device)
    with _k("silu_mul_quant"):
        check(_cabi.load().b200_silu_mul(_ptr(gate_up), 0, _ptr(q), _ptr(s), t, i, qd, _stream()), "b200_silu_mul")
    return q, s


def quantize_per_token(x, quant: str, group_amax=None):
    """x [T, K] bf16 -> (q [T, K], scale [T]).  group_amax [T, n_groups] (from gemm_silu_mul) skips the row scan."""
    _cuda(x, group_amax)
    t, k = x.shape
    assert x.is_contiguous() and x.dtype == torch.bfloat16
    qd = _QDT[quant]
    q = torch.empty((t, k), dtype=_TORCH_Q[qd], device=x.device)
    s = torch.empty((t,), dtype=torch.float32, device=x.device)
    nt = 0
    if group_amax is not None:
        assert group_amax.dtype == torch.float32 and group_amax.is_contiguous() and group_amax.shape[0] == t
        nt = group_amax.shape[1]
    with _k("quantize_per_token"):
        check(_cabi.load().b200_quantize_per_token(_ptr(x), _ptr(q), _ptr(s), _ptr(group_amax), nt, t, k, qd, _stream()),
              "b200_quantize_per_token")
    return q, s


def embedding(tokens, table, out=None):
    _cuda(tokens, table)
    v, h = table.shape
    t = tokens.numel()
    assert tokens.dtype == torch.int64 and table.is_contiguous()
    if out is None:
        out = torch.empty((t, h), dtype=table.dtype, device=table.device)
    with _k("embedding"):
        check(_cabi.load().b200_embedding(_ptr(tokens), _ptr(table), _ptr(out), t, h, v, _stream()), "b200_embedding")
    return out


def argmax(logits, idx_offset: int = 0, want_val: bool = False, out_val=None, out_idx=None):
    _cuda(logits)
    t, v = logits.shape
    assert logits.dtype == torch.float32 and logits.stride(1) == 1
    idx = torch.empty((t,), dtype=torch.int64, device=logits.device) if out_idx is None else out_idx[:t]
    val = (torch.empty((t,), dtype=torch.float32, device=logits.device) if out_val is None else out_val[:t]) if want_val else None
    with _k("argmax"):
        check(_cabi.load().b200_argmax(_ptr(logits), logits.stride(0), _ptr(val), _ptr(idx), t, v, idx_offset, _stream()), "b200_argmax")
    return (val, idx) if want_val else idx


def tp_argmax_merge(peer_vals, peer_idxs, peer_flags, epoch, rank, world, b):
    """Greedy merge of the per-shard (max, global index) pairs over peer memory (tp_comm.TpComm.merge_argmax)."""
    out = torch.empty((b,), dtype=torch.int64, device=epoch.device)
    with _k("tp_argmax_merge"):
        check(_cabi.load().b200_tp_argmax_merge(peer_vals, peer_idxs, peer_flags, _ptr(epoch), rank, world, _ptr(out), b,
                                                _stream()), "b200_tp_argmax_merge")
    return out


SAMPLE_K = 256  # candidates per (row, vocab shard); the reference's _MAX_NEURON_SAMPLING_TOP_K (RUNNER:74)


def topk_candidates(logits, idx_offset: int = 0):
    """logits fp32 [B, V] -> (cand_val [B,256] fp32, cand_idx [B,256] int64) sorted by (value desc, index asc)."""
    _cuda(logits)
    b, v = logits.shape
    assert logits.dtype == torch.float32 and logits.stride(1) == 1
    val = torch.empty((b, SAMPLE_K), dtype=torch.float32, device=logits.device)
    idx = torch.empty((b, SAMPLE_K), dtype=torch.int64, device=logits.device)
    with _k("topk_candidates"):
        check(_cabi.load().b200_topk_candidates(_ptr(logits), logits.stride(0), _ptr(val), _ptr(idx), b, v, idx_offset,
                                                _stream()), "b200_topk_candidates")
    return val, idx


def sample_candidates(cand_val, cand_idx, params, u, out=None):
    """cand_* [B,256] or [groups,B,256]; params fp32 [B,3] = (top_k, top_p, temperature); u fp32 [B] in [0,1)."""
    _cuda(cand_val, cand_idx, params, u)
    if cand_val.dim() == 2:
        cand_val, cand_idx = cand_val.unsqueeze(0), cand_idx.unsqueeze(0)
    g, b, k = cand_val.shape
    assert k == SAMPLE_K and cand_val.is_contiguous() and cand_idx.is_contiguous() and cand_idx.dtype == torch.int64
    assert params.shape == (b, 3) and params.dtype == torch.float32 and params.is_contiguous()
    assert u.shape == (b,) and u.dtype == torch.float32
    if out is None:
        out = torch.empty((b,), dtype=torch.int64, device=cand_val.device)
    with _k("sample_candidates"):
        check(_cabi.load().b200_sample_candidates(_ptr(cand_val), _ptr(cand_idx), g, b, _ptr(params), _ptr(u), _ptr(out),
                                                  _stream()), "b200_sample_candidates")
    return out


def row_lse(logits):
    """logits fp32 [B, V] -> (row max [B], sum exp(x - max) [B]) of the local vocab shard."""
    _cuda(logits)
    b, v = logits.shape
    assert logits.dtype == torch.float32 and logits.stride(1) == 1
    m = torch.empty((b,), dtype=torch.float32, device=logits.device)
    s = torch.empty((b,), dtype=torch.float32, device=logits.device)
    with _k("row_lse"):
        check(_cabi.load().b200_row_lse(_ptr(logits), logits.stride(0), _ptr(m), _ptr(s), b, v, _stream()), "b200_row_lse")
    return m, s


def gather_logprobs(cand_val, cand_idx, sampled, row_max, row_sumexp, num_logprobs: int):
    """cand_* [B,256] or [groups,B,256]; row_max / row_sumexp [B] or [groups,B]; sampled int64 [B].
    -> (ids int32 [B, 1+n], logprobs fp32 [B, 1+n], ranks int32 [B]) in vLLM's LogprobsTensors layout."""
    _cuda(cand_val, cand_idx, sampled, row_max, row_sumexp)
    if cand_val.dim() == 2:
        cand_val, cand_idx, row_max, row_sumexp = cand_val.unsqueeze(0), cand_idx.unsqueeze(0), row_max.unsqueeze(0), row_sumexp.unsqueeze(0)
    g, b, k = cand_val.shape
    assert k == SAMPLE_K and cand_val.is_contiguous() and cand_idx.is_contiguous() and row_max.is_contiguous()
    assert row_max.shape == (g, b) and row_sumexp.shape == (g, b) and row_sumexp.is_contiguous() and sampled.dtype == torch.int64
    ids = torch.empty((b, 1 + num_logprobs), dtype=torch.int32, device=cand_val.device)
    lp = torch.empty((b, 1 + num_logprobs), dtype=torch.float32, device=cand_val.device)
    rk = torch.empty((b,), dtype=torch.int32, device=cand_val.device)
    with _k("gather_logprobs"):
        check(_cabi.load().b200_gather_logprobs(_ptr(cand_val), _ptr(cand_idx), g, b, _ptr(sampled), _ptr(row_max),
                                                _ptr(row_sumexp), num_logprobs, _ptr(ids), _ptr(lp), _ptr(rk), _stream()),
              "b200_gather_logprobs")
    return ids, lp, rk


def allreduce_rmsnorm(peer_bufs, peer_flags, epoch, rank, world, t, h, norm_w, eps, quant=None, want_bf16=False, mc_buf=0):
    """One-shot all-reduce over NVLink peer memory fused with RMSNorm (+ per-token quantisation).
    peer_bufs / peer_flags: ctypes arrays of `world` device pointers (tp_comm.TpComm).  mc_buf != 0: the multicast address
    of the partial -- the sum is then done inside the NVSwitch (b200_allreduce_rmsnorm_nvls).  Returns (x, h_out, h_scale):
    x = the reduced bf16 residual stream; h_out = normed bf16 (quant None) or quantised tensor; h_scale = scales or None."""
    dev = epoch.device
    x = torch.empty((t, h), dtype=torch.bfloat16, device=dev)
    qd = _QDT[quant]
    lib = _cabi.load()
    if mc_buf:
        fn, first, name = lib.b200_allreduce_rmsnorm_nvls, int(mc_buf), "b200_allreduce_rmsnorm_nvls"
    else:
        fn, first, name = lib.b200_allreduce_rmsnorm, peer_bufs, "b200_allreduce_rmsnorm"
    if norm_w is None:
        with _k("allreduce"):
            check(fn(first, peer_flags, _ptr(epoch), rank, world, _ptr(x), 0, 0, 0, 0, t, h, float(eps), BF16, _stream()), name)
        return x, None, None
    if qd == BF16:
        y = torch.empty_like(x)
        with _k("allreduce_rmsnorm"):
            check(fn(first, peer_flags, _ptr(epoch), rank, world, _ptr(x), _ptr(norm_w), _ptr(y), 0, 0, t, h, float(eps), qd,
                     _stream()), name)
        return x, y, None
    q = torch.empty((t, h), dtype=_TORCH_Q[qd], device=dev)
    s = torch.empty((t,), dtype=torch.float32, device=dev)
    y = torch.empty_like(x) if want_bf16 else None
    with _k("allreduce_rmsnorm_quant"):
        check(fn(first, peer_flags, _ptr(epoch), rank, world, _ptr(x), _ptr(norm_w), _ptr(y), _ptr(q), _ptr(s), t, h,
                 float(eps), qd, _stream()), name)
    return x, q, s
