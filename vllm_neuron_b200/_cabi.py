"""ctypes binding of libb200_paged.so (the C ABI declared in include/b200_paged.h).

Torch is only used by the callers for device memory and streams; the library itself sees raw
pointers.  There is NO fallback: if the shared object is missing the import of any op raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200_paged.so")

OK = 0
ABI_VERSION = 8
BF16, FP8_E4M3, INT8, F32 = 0, 1, 2, 3

_p = C.c_void_p
_i = C.c_int
_i64 = C.c_int64
_f = C.c_float
_sz = C.c_size_t

# name -> (restype, argtypes); mirrors include/b200_paged.h one to one
SIGNATURES = {
    "b200_abi_version": (_i, []),
    "b200_last_error": (C.c_char_p, []),
    "b200_device_check": (_i, [_i]),
    "b200_launch_count": (C.c_uint64, []),
    "b200_set_pdl": (_i, [_i]),
    "b200_build_decode_meta": (_i, [_p, _i64, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p]),
    "b200_remask_spec_output": (_i, [_p, _p, _p, _p, _i, _i, _p]),
    "b200_build_prefill_meta": (_i, [_p, _i, _i, _p, _p, _i, _i, _p]),
    "b200_block_table_scatter": (_i, [_p, _i64, _p, _i, _i, _i, _p]),
    "b200_advance_decode": (_i, [_p, _p, _p, _p, _i, _p]),
    "b200_kv_write": (_i, [_p, _p, _i64, _i64, _p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "b200_rope_kv_write": (_i, [_p, _p, _p, _p, _i, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "b200_paged_attn_decode_workspace_bytes": (_sz, [_i, _i, _i, _i, _i, _i]),
    "b200_paged_attn_decode": (
        _i,
        [_p, _i64, _p, _p, _i, _p, _i64, _p, _p, _p, _p, _p, _i, _p, _sz, _i, _i, _i, _i, _i, _i, _i, _f, _p],
    ),
    "b200_paged_attn_decode_static_max_bytes": (_sz, []),
    "b200_paged_attn_decode_static": (
        _i,
        [_p, _i64, _p, _p, _i, _p, _i64, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _i, _f, _p],
    ),
    "b200_paged_attn_prefill": (_i, [_p, _i64, _p, _p, _i, _p, _p, _i, _i, _i, _i, _i, _i, _f, _p]),
    "b200_gemm_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "b200_gemm_decode_plan": (_i, [_i, _i, _i, _i, _i, _i, _i, _p]),
    "b200_gemm_set_trace": (None, [_p]),
    "b200_gemm": (_i, [_p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _sz, _p]),
    "b200_gemm_int8_wo": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p, _sz, _p]),
    "b200_gemm_silu_mul": (_i, [_p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p, _sz, _p]),
    "b200_gemm_rope_kv_write": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p, _sz, _p, _p, _p, _i, _p, _p, _p,
                                     _i, _i, _i, _i, _i, _p]),
    "b200_rmsnorm": (_i, [_p, _p, _p, _p, _p, _i, _i, _f, _i, _p]),
    "b200_silu_mul": (_i, [_p, _p, _p, _p, _i, _i, _i, _p]),
    "b200_quantize_per_token": (_i, [_p, _p, _p, _p, _i, _i, _i, _i, _p]),
    "b200_embedding": (_i, [_p, _p, _p, _i, _i, _i, _p]),
    "b200_argmax": (_i, [_p, _i64, _p, _p, _i, _i, _i64, _p]),
    "b200_topk_candidates": (_i, [_p, _i64, _p, _p, _i, _i, _i64, _p]),
    "b200_sample_candidates": (_i, [_p, _p, _i, _i, _p, _p, _p, _p]),
    "b200_row_lse": (_i, [_p, _i64, _p, _p, _i, _i, _p]),
    "b200_gather_logprobs": (_i, [_p, _p, _i, _i, _p, _p, _p, _i, _p, _p, _p, _p]),
    "b200_allreduce_rmsnorm": (_i, [_p, _p, _p, _i, _i, _p, _p, _p, _p, _p, _i, _i, _f, _i, _p]),
    "b200_gemm_silu_mul_quant": (_i, [_p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p]),
    "b200_tp_argmax_merge": (_i, [_p, _p, _p, _p, _i, _i, _p, _i, _p]),
    "b200_allreduce_rmsnorm_nvls": (_i, [_p, _p, _p, _i, _i, _p, _p, _p, _p, _p, _i, _i, _f, _i, _p]),
}

_lib = None


class B200Error(RuntimeError):
    pass


def load():
    """Load the shared object once.  Raises (loudly) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200Error(
            f"{LIB_PATH} is missing: build it with `python -m vllm_neuron_b200.build` "
            "(there is no CPU or library fallback for the hot path)"
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    if lib.b200_abi_version() != ABI_VERSION:
        raise B200Error(f"ABI version mismatch: library {lib.b200_abi_version()} != binding {ABI_VERSION}")
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != OK:
        msg = load().b200_last_error().decode(errors="replace")
        raise B200Error(f"{what} failed (rc={rc}): {msg}")
