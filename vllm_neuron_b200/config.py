"""Model-shape and block-KV configuration for the B200 path.

`B200Config.from_vllm` is the counterpart of the reference's `_get_default_neuron_config` /
`_handle_pa_num_blocks` / `_validate_neuron_config` / `_get_neuron_config_after_override`
(vllm_neuron/worker/neuronx_distributed_model_loader.py:725-900): it maps vLLM's configs plus
`additional_config["override_neuron_config"]` to the block-KV keys the hot path needs, with the same
key names, the same block-count rules and the same errors.
"""
from __future__ import annotations

from dataclasses import dataclass, field, replace
from math import ceil

# Neuron-compiler-only keys: accepted for drop-in compatibility, ignored on B200.
IGNORED_NEURON_KEYS = frozenset(
    {
        "skip_warmup", "enable_bucketing", "context_encoding_buckets", "token_generation_buckets",
        "save_sharded_checkpoint", "async_mode", "sequence_parallel_enabled", "fused_qkv", "logical_nc_config",
        "cc_pipeline_tiling_factor", "attn_kernel_enabled", "qkv_kernel_enabled", "mlp_kernel_enabled",
        "attn_tkg_nki_kernel_enabled", "attn_block_tkg_nki_kernel_enabled", "on_device_sampling_config",
        "padding_side", "is_continuous_batching", "ctx_batch_size", "torch_dtype", "lora_config",
    }
)


@dataclass
class ModelShape:
    name: str
    hidden: int
    layers: int
    num_q_heads: int
    num_kv_heads: int
    head_dim: int
    intermediate: int
    vocab: int
    rope_theta: float = 10000.0
    rope_scaling: dict | None = None
    rms_eps: float = 1e-5
    qkv_bias: bool = False

    @staticmethod
    def from_hf(hf) -> "ModelShape":
        """LOADER:612-631 (_get_model_configs) equivalent for Llama/Qwen2-family HF configs."""
        heads = hf.num_attention_heads
        head_dim = getattr(hf, "head_dim", None) or hf.hidden_size // heads
        scaling = getattr(hf, "rope_scaling", None)
        if scaling is None:
            rp = getattr(hf, "rope_parameters", None)
            if isinstance(rp, dict) and rp.get("rope_type", "default") != "default":
                scaling = rp
        theta = getattr(hf, "rope_theta", None)
        if theta is None:
            theta = (getattr(hf, "rope_parameters", None) or {}).get("rope_theta", 10000.0)
        arch = (getattr(hf, "architectures", None) or [""])[0]
        return ModelShape(
            name=getattr(hf, "_name_or_path", "") or arch,
            hidden=hf.hidden_size,
            layers=hf.num_hidden_layers,
            num_q_heads=heads,
            num_kv_heads=getattr(hf, "num_key_value_heads", heads) or heads,
            head_dim=head_dim,
            intermediate=hf.intermediate_size,
            vocab=hf.vocab_size,
            rope_theta=float(theta),
            rope_scaling=dict(scaling) if scaling else None,
            rms_eps=getattr(hf, "rms_norm_eps", 1e-5),
            qkv_bias=bool(getattr(hf, "attention_bias", False)) or "Qwen2" in arch,
        )


_LLAMA3_SCALING = {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                   "original_max_position_embeddings": 8192}

PRESETS = {
    # BASELINE.json configs[0]
    "tinyllama-1.1b": ModelShape("TinyLlama-1.1B", 2048, 22, 32, 4, 64, 5632, 32000, 10000.0, None, 1e-5),
    # configs[1], configs[2]
    "llama-3.1-8b": ModelShape("Llama-3.1-8B", 4096, 32, 32, 8, 128, 14336, 128256, 500000.0, _LLAMA3_SCALING, 1e-5),
    # configs[3]
    "qwen2.5-32b": ModelShape("Qwen2.5-32B", 5120, 64, 40, 8, 128, 27648, 152064, 1000000.0, None, 1e-6, True),
    # configs[4]
    "llama-3.3-70b": ModelShape("Llama-3.3-70B", 8192, 80, 64, 8, 128, 28672, 128256, 500000.0, _LLAMA3_SCALING, 1e-5),
    # parity at the headline WIDTH: Llama-3.1-8B's per-layer dimensions (and vocabulary), two layers deep, so the CPU
    # oracle still finishes in minutes (tests/test_gpu_model.py::test_fp8_parity_at_headline_width)
    "llama-3.1-8b-2l": ModelShape("Llama-3.1-8B-2L", 4096, 2, 32, 8, 128, 14336, 128256, 500000.0, _LLAMA3_SCALING, 1e-5),
    # unit-test size
    "tiny-test": ModelShape("tiny-test", 256, 2, 8, 2, 64, 512, 1024, 10000.0, None, 1e-5),
    # unit-test size with Qwen2's traits: 5 query heads per kv head, qkv bias, vocab not a multiple of the GEMM tile
    "tiny-qwen": ModelShape("tiny-qwen", 640, 2, 10, 2, 64, 1152, 1000, 1000000.0, None, 1e-6, True),
}


@dataclass
class B200Config:
    """The neuron_config keys that matter on the hot path, under their reference names."""

    tp_degree: int = 1
    batch_size: int = 32          # max_num_seqs
    seq_len: int = 2048           # max_model_len
    max_context_length: int = 2048
    pa_num_blocks: int = 0
    pa_block_size: int = 32
    is_block_kv_layout: bool = True
    is_prefix_caching: bool = True
    quantized: bool = False
    quantization_type: str = "per_channel_symmetric"
    quantization_dtype: str = "int8"     # "int8" | "f8e4m3"
    weight_only: bool = False
    modules_to_not_convert: tuple = ()
    quantized_checkpoints_path: str | None = None   # LOADER:890-891: where the once-quantised state dict lives
    ignored: dict = field(default_factory=dict)

    @property
    def quant_mode(self) -> str:
        """'bf16' | 'fp8' | 'int8' | 'int8_wo' as used by model.py."""
        if not self.quantized:
            return "bf16"
        if self.quantization_dtype in ("f8e4m3", "fp8", "float8_e4m3fn", "f8e4m3fn"):
            return "fp8"
        if self.quantization_dtype == "int8":
            return "int8_wo" if self.weight_only else "int8"
        raise ValueError(f"unsupported quantization_dtype {self.quantization_dtype!r}")

    @property
    def max_blocks_per_seq(self) -> int:
        return self.seq_len // self.pa_block_size  # floor, as RUNNER:749,814


def default_num_blocks(max_model_len: int, block_size: int, max_num_seqs: int, num_gpu_blocks_override) -> int:
    """LOADER:741-745 (floor-then-ceil kept as in the reference)."""
    n = ceil(max_model_len // block_size) * max_num_seqs
    return num_gpu_blocks_override if num_gpu_blocks_override is not None else n


def handle_pa_num_blocks(num_gpu_blocks_override, cfg: dict, override: dict | None) -> dict:
    """LOADER:796-831 -- keep vLLM's (+1 null block) count and pa_num_blocks consistent."""
    explicit = bool(override) and "pa_num_blocks" in override
    if num_gpu_blocks_override is not None:
        pa = cfg.get("pa_num_blocks")
        original_user_override = num_gpu_blocks_override - 1
        if explicit:
            if pa == original_user_override:
                cfg["pa_num_blocks"] = num_gpu_blocks_override
            else:
                raise ValueError(
                    f"pa_num_blocks ({pa}) must match your --num-gpu-blocks-override intent({original_user_override}) "
                    "to ensure vLLM and the B200 KV pool have consistent block counts. "
                )
    elif explicit:
        raise ValueError(
            f"When setting pa_num_blocks ({cfg.get('pa_num_blocks')}) in override_neuron_config, you must also set "
            "--num-gpu-blocks-override to the same value to ensure consistent block counts."
        )
    return cfg


def validate_config(enable_prefix_caching: bool, max_model_len: int, block_size: int, max_num_seqs: int,
                    num_gpu_blocks_override, cfg: dict) -> dict:
    """LOADER:834-867."""
    if enable_prefix_caching:
        assert cfg.get("is_prefix_caching", False)
        assert cfg.get("is_block_kv_layout", False)
    if cfg.get("is_block_kv_layout"):
        min_blocks_required = ceil(max_model_len / block_size) * max_num_seqs
        effective = num_gpu_blocks_override - 1 if num_gpu_blocks_override is not None else cfg.get("pa_num_blocks")
        assert effective >= min_blocks_required, (
            f"At least {min_blocks_required} blocks are required for max_model_len {max_model_len}, but only "
            f"{effective} blocks are available (user-intended blocks, excluding the +1 for null block)"
        )
    return cfg


def build_config(*, tensor_parallel_size: int, max_num_seqs: int, max_model_len: int, block_size: int,
                 enable_prefix_caching: bool, num_gpu_blocks_override, override: dict | None) -> B200Config:
    """_get_default_neuron_config -> override merge -> _handle_pa_num_blocks -> _validate_neuron_config."""
    override = dict(override or {})
    cfg = {
        "tp_degree": tensor_parallel_size,
        "batch_size": max_num_seqs,
        "max_context_length": max_model_len,
        "seq_len": max_model_len,
        "quantized": False,
        "pa_num_blocks": default_num_blocks(max_model_len, block_size, max_num_seqs, num_gpu_blocks_override),
        "pa_block_size": block_size,
        # B200 always runs the block layout; the reference enables it for prefix caching / chunked prefill
        "is_block_kv_layout": True,
        "is_prefix_caching": bool(enable_prefix_caching),
    }
    merged = dict(override)
    if "quantized" in merged:  # LOADER:885-898
        cfg.update(
            quantized=merged.pop("quantized", False),
            quantized_checkpoints_path=merged.pop("quantized_checkpoints_path", None),
            quantization_type=merged.pop("quantization_type", "per_tensor_symmetric"),
            quantization_dtype=merged.pop("quantization_dtype", "int8"),
        )
    cfg.update(merged)
    if cfg.get("quantized") and cfg.get("quantization_type", "per_channel_symmetric") not in (
            "per_channel_symmetric", "per_tensor_symmetric"):
        raise ValueError(f"quantization_type {cfg['quantization_type']!r} is not supported on the B200 path "
                         "(per_channel_symmetric or per_tensor_symmetric)")
    cfg = handle_pa_num_blocks(num_gpu_blocks_override, cfg, override)
    cfg = validate_config(enable_prefix_caching, max_model_len, block_size, max_num_seqs, num_gpu_blocks_override, cfg)
    known = {f for f in B200Config.__dataclass_fields__ if f != "ignored"}
    kept = {k: v for k, v in cfg.items() if k in known}
    ignored = {k: v for k, v in cfg.items() if k not in known}
    unknown = [k for k in ignored if k not in IGNORED_NEURON_KEYS and not k.endswith("_buckets") and not k.endswith("_config")]
    if unknown:
        raise ValueError(f"unknown override_neuron_config keys for the B200 path: {sorted(unknown)}")
    if "modules_to_not_convert" in kept and kept["modules_to_not_convert"] is not None:
        kept["modules_to_not_convert"] = tuple(kept["modules_to_not_convert"])
    return B200Config(ignored=ignored, **kept)
