"""Checkpoint loading for the B200 path: HF safetensors -> the weight dictionary B200CausalLM consumes.

Reference: vllm_neuron/worker/neuronx_distributed_model_loader.py
  * model path / HF config lookup             :612-660   (_get_model_configs, get_neuron_model)
  * quantise once, save, reload               :234-241   (save_quantized_state_dict + quantized_checkpoints_path)
  * quantisation keys                         :885-898   (quantized, quantized_checkpoints_path, quantization_type/_dtype)
  * modules_to_not_convert                    README / test/utils/server/server.py:307-333 (passed through to NxDI)
The reference hands a model directory to NxDI, which reads the HF state dict, fuses q/k/v, shards and (optionally)
quantises it.  Here the same steps are explicit and lazy, one decoder layer at a time, so that an 8B-70B checkpoint never
needs a full host copy:  HF tensor names -> {embed, lm_head, final_norm, layers[i]{input_norm, post_norm, wqkv, wo, wgu,
wdown[, bqkv]}} (un-sharded, rows = output channels); B200CausalLM quantises per output channel and takes its TP shard.

A quantised checkpoint (`quantized_checkpoints_path`) stores for every linear `<name>.weight` in the GEMM dtype
(float8_e4m3fn / int8) and `<name>.scale` ([N] fp32, per output channel); a layer read from it is handed on as a
(weight, scale) pair and is not quantised again.
"""
from __future__ import annotations

import glob
import json
import os

import torch

from .config import ModelShape

_LAYER = "model.layers.{}."
_HF_LINEARS = ("self_attn.q_proj", "self_attn.k_proj", "self_attn.v_proj", "self_attn.o_proj",
               "mlp.gate_proj", "mlp.up_proj", "mlp.down_proj")


class CheckpointError(RuntimeError):
    pass


def _index_safetensors(path: str) -> dict[str, str]:
    """tensor name -> file, from model.safetensors.index.json when present, else by opening every *.safetensors."""
    from safetensors import safe_open

    idx = os.path.join(path, "model.safetensors.index.json")
    if os.path.exists(idx):
        with open(idx) as f:
            return {k: os.path.join(path, v) for k, v in json.load(f)["weight_map"].items()}
    files = sorted(glob.glob(os.path.join(path, "*.safetensors")))
    if not files:
        raise CheckpointError(f"no *.safetensors checkpoint under {path!r} (torch .bin checkpoints are not supported)")
    out: dict[str, str] = {}
    for fn in files:
        with safe_open(fn, framework="pt") as f:
            for k in f.keys():
                out[k] = fn
    return out


class HFCheckpointWeights:
    """Lazy view of an HF Llama/Qwen2-family checkpoint directory in B200CausalLM's weight-dictionary protocol
    (`w["embed"]`, `w["lm_head"]`, `w["final_norm"]`, `w["layers"][i]`, `w.drop(i)`)."""

    def __init__(self, path: str, shape: ModelShape, *, tie_word_embeddings: bool = False):
        self.path, self.shape, self.tie = path, shape, tie_word_embeddings
        self.index = _index_safetensors(path)
        self._handles: dict = {}
        self._cache: dict = {}
        self.layers = _Layers(self)
        self._check()

    # ---- raw access
    def _get(self, name: str) -> torch.Tensor:
        from safetensors import safe_open

        fn = self.index.get(name)
        if fn is None:
            raise CheckpointError(f"tensor {name!r} is missing from the checkpoint at {self.path!r}")
        if fn not in self._handles:
            self._handles[fn] = safe_open(fn, framework="pt")
        return self._handles[fn].get_tensor(name)

    def _has(self, name: str) -> bool:
        return name in self.index

    def _linear(self, name: str, n: int, k: int):
        """bf16 [N,K] master weight, or (q[N,K], scale[N]) when the checkpoint is already quantised."""
        w = self._get(name + ".weight")
        if tuple(w.shape) != (n, k):
            raise CheckpointError(f"{name}.weight has shape {tuple(w.shape)}, the model config says {(n, k)}")
        if w.dtype in (torch.float8_e4m3fn, torch.int8):
            s = self._get(name + ".scale").float().reshape(-1)
            if s.numel() != n:
                raise CheckpointError(f"{name}.scale must hold one fp32 scale per output channel ({n}), got {s.numel()}")
            return (w, s)
        return w.to(torch.bfloat16)

    def _check(self):
        sh = self.shape
        need = ["model.embed_tokens.weight", "model.norm.weight"]
        for li in (0, sh.layers - 1):
            need += [_LAYER.format(li) + m + ".weight" for m in _HF_LINEARS]
        missing = [n for n in need if not self._has(n)]
        if missing:
            raise CheckpointError(f"checkpoint at {self.path!r} lacks {missing[:4]}{'...' if len(missing) > 4 else ''}")
        if self._has(_LAYER.format(sh.layers) + "self_attn.q_proj.weight"):
            raise CheckpointError(f"checkpoint has more than num_hidden_layers={sh.layers} decoder layers")

    # ---- weight-dictionary protocol
    def __getitem__(self, k):
        sh = self.shape
        if k == "layers":
            return self.layers
        if k == "embed":
            return self._get("model.embed_tokens.weight").to(torch.bfloat16)
        if k == "final_norm":
            return self._get("model.norm.weight").to(torch.bfloat16)
        if k == "lm_head":
            if self._has("lm_head.weight") and not self.tie:
                return self._linear("lm_head", sh.vocab, sh.hidden)
            return self._get("model.embed_tokens.weight").to(torch.bfloat16)  # tie_word_embeddings
        raise KeyError(k)

    def layer(self, li: int) -> dict:
        if li in self._cache:
            return self._cache[li]
        sh, p = self.shape, _LAYER.format(li)
        qd, kd = sh.num_q_heads * sh.head_dim, sh.num_kv_heads * sh.head_dim
        q = self._linear(p + "self_attn.q_proj", qd, sh.hidden)
        k = self._linear(p + "self_attn.k_proj", kd, sh.hidden)
        v = self._linear(p + "self_attn.v_proj", kd, sh.hidden)
        g = self._linear(p + "mlp.gate_proj", sh.intermediate, sh.hidden)
        u = self._linear(p + "mlp.up_proj", sh.intermediate, sh.hidden)
        lw = {
            "input_norm": self._get(p + "input_layernorm.weight").to(torch.bfloat16),
            "post_norm": self._get(p + "post_attention_layernorm.weight").to(torch.bfloat16),
            "wqkv": _cat_rows(q, k, v),
            "wo": self._linear(p + "self_attn.o_proj", sh.hidden, qd),
            "wgu": _cat_rows(g, u),
            "wdown": self._linear(p + "mlp.down_proj", sh.hidden, sh.intermediate),
        }
        if sh.qkv_bias:
            lw["bqkv"] = torch.cat([self._get(p + f"self_attn.{n}_proj.bias").to(torch.bfloat16) for n in "qkv"])
        self._cache[li] = lw
        return lw

    def drop(self, li: int) -> None:
        self._cache.pop(li, None)


class _Layers:
    def __init__(self, parent):
        self.parent = parent

    def __getitem__(self, li):
        return self.parent.layer(li)


def _cat_rows(*parts):
    """Stack linears along the output-channel axis; per-channel scales stack the same way (they belong to rows)."""
    if isinstance(parts[0], tuple):
        assert all(isinstance(p, tuple) for p in parts), "q/k/v (gate/up) must be quantised alike"
        return (torch.cat([p[0].view(torch.uint8) for p in parts]).view(parts[0][0].dtype),
                torch.cat([p[1] for p in parts]))
    return torch.cat(parts, dim=0)


def skip_quantization(module_name: str, modules_to_not_convert) -> bool:
    """True when `module_name` (an HF module path such as "lm_head" or "model.layers.3.mlp.down_proj") matches an entry
    of modules_to_not_convert: exact name, dotted-path suffix or prefix (how HF quantisers and NxDI read the key)."""
    for m in modules_to_not_convert or ():
        if module_name == m or module_name.endswith("." + m) or module_name.startswith(m + "."):
            return True
    return False


def check_modules_to_not_convert(modules_to_not_convert, shape: ModelShape) -> bool:
    """The fused decode chain hands quantised activations from kernel to kernel, so decoder-layer linears are converted
    all or none; the one module that can stay in bf16 on its own is lm_head.  Returns quant_lm_head."""
    mods = tuple(modules_to_not_convert or ())
    other = [m for m in mods if not skip_quantization("lm_head", [m])]
    if other:
        raise ValueError(
            f"modules_to_not_convert={other}: only 'lm_head' can be left unquantised on the B200 path "
            "(decoder-layer GEMMs consume the previous kernel's quantised activations)")
    return not mods


def save_quantized_checkpoint(src: "HFCheckpointWeights", out_dir: str, kind: str, modules_to_not_convert=(),
                              quant_type: str = "per_channel_symmetric") -> str:
    """Quantise an HF checkpoint once (symmetric, per output channel or per tensor; kind 'fp8' | 'int8') and write it as one safetensors
    file per decoder layer + model.safetensors.index.json.  Counterpart of save_quantized_state_dict (LOADER:237-239);
    the result is what `quantized_checkpoints_path` points at."""
    from safetensors.torch import save_file

    from .model import QUANT_TYPES, quantize_weight

    assert kind in ("fp8", "int8")
    if quant_type not in QUANT_TYPES:
        raise ValueError(f"quantization_type {quant_type!r} is not supported on the B200 path (one of {QUANT_TYPES})")
    per_tensor = quant_type == "per_tensor_symmetric"
    os.makedirs(out_dir, exist_ok=True)
    weight_map: dict[str, str] = {}

    def put(tensors: dict, fn: str):
        save_file({k: v.contiguous() for k, v in tensors.items()}, os.path.join(out_dir, fn))
        weight_map.update({k: fn for k in tensors})

    def conv(name: str, out: dict):
        w = src._get(name + ".weight")
        if skip_quantization(name, modules_to_not_convert):
            out[name + ".weight"] = w.to(torch.bfloat16)
        else:
            q, s = quantize_weight(w.to(torch.bfloat16), kind, [(0, w.shape[0])] if per_tensor else None)
            out[name + ".weight"], out[name + ".scale"] = q, s

    top = {"model.embed_tokens.weight": src._get("model.embed_tokens.weight").to(torch.bfloat16),
           "model.norm.weight": src._get("model.norm.weight").to(torch.bfloat16)}
    if src._has("lm_head.weight") and not src.tie:
        conv("lm_head", top)
    put(top, "model-top.safetensors")
    for li in range(src.shape.layers):
        p, out = _LAYER.format(li), {}
        for m in _HF_LINEARS:
            conv(p + m, out)
        for n in ("input_layernorm.weight", "post_attention_layernorm.weight"):
            out[p + n] = src._get(p + n).to(torch.bfloat16)
        if src.shape.qkv_bias:
            for n in "qkv":
                out[p + f"self_attn.{n}_proj.bias"] = src._get(p + f"self_attn.{n}_proj.bias").to(torch.bfloat16)
        put(out, f"model-layer-{li:05d}.safetensors")
    with open(os.path.join(out_dir, "model.safetensors.index.json"), "w") as f:
        json.dump({"metadata": {"quantization_dtype": kind, "quantization_type": quant_type},
                   "weight_map": weight_map}, f)
    return out_dir
