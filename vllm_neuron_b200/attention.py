"""The single attention backend of the B200 platform: block-KV paged attention on libb200_paged.so.

North-star item (1) "platform and attention-backend selection": there is nothing to select -- the platform names this
class for every request (`B200Platform.get_attn_backend_cls`), and the model (model.py) calls the two entry points
below directly.  No Triton, no FlashInfer/FlashAttention dispatch, no CPU fallback.
"""
from __future__ import annotations

from . import ops


class B200PagedAttentionBackend:
    accept_output_buffer = True
    supported_block_sizes = (32,)       # pa_block_size (reference README.md:37)
    supported_head_sizes = (64, 128)
    kv_cache_dtype = "bfloat16"

    @staticmethod
    def get_name() -> str:
        return "B200_PAGED"

    @staticmethod
    def get_supported_head_sizes():
        return list(B200PagedAttentionBackend.supported_head_sizes)

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int):
        """[2 (K,V)][num_blocks][Hkv][block_size][D]: one (block, kv head) tile is a contiguous 32 x D slab (TMA box)."""
        return (2, num_blocks, num_kv_heads, block_size, head_size)

    decode = staticmethod(ops.paged_attn_decode)      # q_len = 1, flash-decoding split along the context
    prefill = staticmethod(ops.paged_attn_prefill)    # batch 1, causal over (cached prefix || new tokens)
    write_kv = staticmethod(ops.rope_kv_write)        # fused RoPE + scatter by slot_mapping
