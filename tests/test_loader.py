"""CPU tests of checkpoint loading (LOADER:234-241, 885-898): HF safetensors -> the model's weight
dictionary, quantise-once-and-reload (`quantized_checkpoints_path`), `modules_to_not_convert`, and the runner's
load_model wiring.  No kernels run here; the GPU leg (same tokens from checkpoint and from the in-memory dictionary) is
tests/test_gpu_model.py::test_checkpoint_load_matches_in_memory_weights."""
import os

import pytest
import torch

from parity_util import export_hf_checkpoint, make_vllm_config
from vllm_neuron_b200.config import PRESETS
from vllm_neuron_b200.loader import (CheckpointError, HFCheckpointWeights, check_modules_to_not_convert,
                                     save_quantized_checkpoint, skip_quantization)
from vllm_neuron_b200.model import _Linear, quantize_weight, random_init_full_weights, tp_shard_plan


def _same(a, b):
    return a.dtype == b.dtype and a.shape == b.shape and torch.equal(a.view(torch.uint8), b.view(torch.uint8))


@pytest.mark.parametrize("preset,shards", [("tiny-test", 1), ("tiny-qwen", 3)])
def test_hf_checkpoint_round_trip(tmp_path, preset, shards):
    shape = PRESETS[preset]
    w = random_init_full_weights(shape, seed=3)
    src = HFCheckpointWeights(export_hf_checkpoint(w, shape, str(tmp_path), shards=shards), shape)
    for k in ("embed", "lm_head", "final_norm"):
        assert _same(src[k], w[k]), k
    for li in range(shape.layers):
        got = src["layers"][li]
        assert set(got) == set(w["layers"][li])
        for k, v in w["layers"][li].items():
            assert _same(got[k], v), (li, k)
        src.drop(li)
    assert not src._cache


def test_tied_embeddings_and_errors(tmp_path):
    shape = PRESETS["tiny-test"]
    w = random_init_full_weights(shape, seed=4)
    d = export_hf_checkpoint(w, shape, str(tmp_path / "tied"), shards=1, lm_head=False)
    assert _same(HFCheckpointWeights(d, shape)["lm_head"], w["embed"])
    with pytest.raises(CheckpointError, match="no \\*.safetensors"):
        HFCheckpointWeights(str(tmp_path / "nothing"), shape)
    import dataclasses

    with pytest.raises(CheckpointError, match="lacks"):
        HFCheckpointWeights(d, dataclasses.replace(shape, layers=3))
    with pytest.raises(CheckpointError, match="more than num_hidden_layers"):
        HFCheckpointWeights(d, dataclasses.replace(shape, layers=1))
    with pytest.raises(CheckpointError, match="shape"):
        HFCheckpointWeights(d, dataclasses.replace(shape, intermediate=256))["layers"][0]


@pytest.mark.parametrize("kind,mode", [("fp8", "fp8"), ("int8", "int8"), ("int8", "int8_wo")])
def test_quantise_once_reload_is_bit_identical(tmp_path, kind, mode):
    """q/k/v quantised one by one and stacked == the fused wqkv quantised whole (scales are per output row), and a
    _Linear built from the reloaded pair == one built from the bf16 master, shard by shard."""
    shape = PRESETS["tiny-qwen"]
    w = random_init_full_weights(shape, seed=5)
    src = HFCheckpointWeights(export_hf_checkpoint(w, shape, str(tmp_path / "hf")), shape)
    qdir = save_quantized_checkpoint(src, str(tmp_path / "q"), kind)
    assert os.path.exists(os.path.join(qdir, "model.safetensors.index.json"))
    q = HFCheckpointWeights(qdir, shape)
    for li in range(shape.layers):
        for name in ("wqkv", "wo", "wgu", "wdown"):
            wq, ws = quantize_weight(w["layers"][li][name], kind)
            got = q["layers"][li][name]
            assert isinstance(got, tuple) and _same(got[0], wq) and _same(got[1], ws), (li, name)
        assert _same(q["layers"][li]["bqkv"], w["layers"][li]["bqkv"])
    for rank in range(2):
        plan = tp_shard_plan(shape, rank, 2)
        a = _Linear(w["layers"][0]["wqkv"], mode, "cpu", row_slices=plan["qkv_rows"], bias=w["layers"][0]["bqkv"])
        b = _Linear(q["layers"][0]["wqkv"], mode, "cpu", row_slices=plan["qkv_rows"], bias=q["layers"][0]["bqkv"])
        assert _same(a.w, b.w) and _same(a.scale, b.scale) and _same(a.bias, b.bias)
        a = _Linear(w["layers"][1]["wdown"], mode, "cpu", col_slice=plan["down_cols"])
        b = _Linear(q["layers"][1]["wdown"], mode, "cpu", col_slice=plan["down_cols"])
        assert _same(a.w, b.w) and _same(a.scale, b.scale)
    with pytest.raises(ValueError, match="quantised checkpoint holds"):
        _Linear(q["layers"][0]["wo"], "fp8" if kind == "int8" else "int8", "cpu")


def test_modules_to_not_convert(tmp_path):
    shape = PRESETS["tiny-test"]
    assert skip_quantization("lm_head", ["lm_head"]) and skip_quantization("model.layers.0.mlp.down_proj", ["down_proj"])
    assert skip_quantization("model.layers.0.mlp.down_proj", ["model.layers.0"]) and not skip_quantization("lm_head", ["head"])
    assert check_modules_to_not_convert((), shape) is True and check_modules_to_not_convert(("lm_head",), shape) is False
    with pytest.raises(ValueError, match="only 'lm_head'"):
        check_modules_to_not_convert(("lm_head", "model.layers.0.mlp.down_proj"), shape)
    w = random_init_full_weights(shape, seed=6)
    src = HFCheckpointWeights(export_hf_checkpoint(w, shape, str(tmp_path / "hf")), shape)
    q = HFCheckpointWeights(save_quantized_checkpoint(src, str(tmp_path / "q"), "fp8", ["lm_head"]), shape)
    assert _same(q["lm_head"], w["lm_head"]) and isinstance(q["layers"][0]["wo"], tuple)


def test_runner_load_model_opens_checkpoint_and_quantises_once(tmp_path):
    from vllm_neuron_b200.worker.b200_model_runner import B200ModelRunner

    shape = PRESETS["tiny-test"]
    w = random_init_full_weights(shape, seed=7)
    hf = export_hf_checkpoint(w, shape, str(tmp_path / "hf"))
    qdir = str(tmp_path / "quantised")
    override = {"quantized": True, "quantization_dtype": "f8e4m3", "quantization_type": "per_channel_symmetric",
                "quantized_checkpoints_path": qdir, "modules_to_not_convert": ["lm_head"]}
    cfg = make_vllm_config(shape, max_num_seqs=2, max_model_len=128, num_gpu_blocks_override=9, override=override)
    cfg.model_config.model = hf
    cfg.load_config.load_format = "auto"
    r = B200ModelRunner(cfg, torch.device("cpu"))
    r.load_model()
    assert r.b200_config.quantized_checkpoints_path == qdir and r._quant_lm_head is False
    assert r._weights.path == qdir and isinstance(r._weights["layers"][1]["wgu"], tuple)
    stamp = os.path.getmtime(os.path.join(qdir, "model.safetensors.index.json"))
    r2 = B200ModelRunner(cfg, torch.device("cpu"))
    r2.load_model()  # second start: read back, not re-quantised
    assert os.path.getmtime(os.path.join(qdir, "model.safetensors.index.json")) == stamp
    cfg.load_config.load_format = "dummy"
    r3 = B200ModelRunner(cfg, torch.device("cpu"))
    r3.load_model()
    assert r3._weights is None


def test_per_tensor_symmetric_scales_follow_the_original_tensors():
    """quantization_type per_tensor_symmetric (the reference's default, LOADER:893-895): one scale per ORIGINAL tensor
    (q, k, v and gate, up are separate HF tensors), identical between the product quantiser and the oracle's."""
    import torch

    from oracle import quant_ref
    from vllm_neuron_b200.model import quantize_weight

    w = (torch.randn(96, 64, generator=torch.Generator().manual_seed(2)) * 0.05).to(torch.bfloat16)
    rows = [(0, 64), (64, 80), (80, 96)]
    for kind in ("fp8", "int8"):
        q, s = quantize_weight(w, kind, rows)
        for a, b in rows:
            amax = w[a:b].float().abs().max()
            assert torch.all(s[a:b] == s[a]) and abs(s[a].item() - amax.item() / (448.0 if kind == "fp8" else 127.0)) < 1e-9
        q2, s2 = quant_ref.quantize_weight_per_channel(w, kind, rows)
        assert torch.equal(s, s2) and torch.equal(q.view(torch.uint8), q2.view(torch.uint8))
        qc, sc = quantize_weight(w, kind)
        assert not torch.all(sc[:64] == sc[0])  # per channel differs


def test_quantization_type_default_and_validation():
    import pytest

    from vllm_neuron_b200.config import build_config

    kw = dict(tensor_parallel_size=1, max_num_seqs=2, max_model_len=64, block_size=32, enable_prefix_caching=True,
              num_gpu_blocks_override=None)
    cfg = build_config(override={"quantized": True, "quantization_dtype": "f8e4m3"}, **kw)
    assert cfg.quantization_type == "per_tensor_symmetric"  # the reference's default
    with pytest.raises(ValueError):
        build_config(override={"quantized": True, "quantization_type": "blockwise"}, **kw)
