"""CPU: the C-ABI library builds, loads, and exports exactly what include/b200_paged.h declares."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from vllm_neuron_b200 import build, _cabi

    build.build()
    return _cabi.load()


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "b200_paged.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree(lib):
    from vllm_neuron_b200 import _cabi

    names = _header_symbols()
    assert len(names) >= 18
    assert sorted(_cabi.SIGNATURES) == names
    for n in names:
        assert hasattr(lib, n), n


def test_abi_version(lib):
    from vllm_neuron_b200 import _cabi
    assert lib.b200_abi_version() == _cabi.ABI_VERSION == 8


def test_no_device_is_an_error_not_a_fallback(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    rc = lib.b200_device_check(0)
    assert rc == -2  # B200_ERR_NO_DEVICE
    assert b"no CUDA device" in lib.b200_last_error()


def test_argument_validation_without_gpu(lib):
    # null pointers / bad shapes are rejected before any launch
    assert lib.b200_gemm(None, None, None, None, None, None, None, 1, 1, 16, 0, 0, 0, None, 0, None) == -1
    assert lib.b200_paged_attn_decode(None, 0, None, None, 0, None, 0, None, None, None, None, None, 0, None, 0, 1, 8, 8, 128, 32, 4, 0, 1.0, None) == -1
    # balanced split: at most ceil(max_blocks / min_quota) + 1 partials per (sequence, head)
    assert lib.b200_paged_attn_decode_workspace_bytes(64, 32, 8, 128, 64, 8) == 64 * 32 * 9 * 130 * 4 + 64 * 8 * 4  # + part counts per (sequence, kv head)


def test_ops_refuse_cpu_tensors():
    import torch

    from vllm_neuron_b200 import _cabi, ops

    x = torch.zeros(2, 64, dtype=torch.bfloat16)
    with pytest.raises(_cabi.B200Error):
        ops.rmsnorm(x, torch.ones(64, dtype=torch.bfloat16), 1e-5)


def test_product_does_not_import_oracle():
    import subprocess
    import sys

    out = subprocess.run(
        ["grep", "-rnE", r"^\s*(from|import)\s+oracle", os.path.join(ROOT, "vllm_neuron_b200")], capture_output=True, text=True
    )
    assert out.stdout == "", out.stdout
