"""Oracle (test infrastructure): integer index / config semantics of the block-KV path.

Each function restates one piece of the reference's Python and cites it.  Paths
are relative to /root/reference; RUNNER = vllm_neuron/worker/
neuronx_distributed_model_runner.py, LOADER = vllm_neuron/worker/
neuronx_distributed_model_loader.py, PLAT = vllm_neuron/platform.py.

Parity: PINNED against golden vectors generated from the reference's own code
(tests/golden/gen_from_reference.py -> tests/golden/index_golden.json).
Pure Python / numpy on purpose: the sizes are tiny and the loops mirror the
reference's loops.
"""
from __future__ import annotations

from math import ceil

import numpy as np

SLOT_MAPPING_PAD = -1  # RUNNER:80
BLOCK_TABLE_PAD = 0  # RUNNER:81
MAX_SAMPLING_TOP_K = 256  # RUNNER:74


def last_token_position(prompt_len: int, num_output_tokens: int) -> int:
    """RUNNER:156-175 -- 0-indexed position of the last processed token."""
    return prompt_len + num_output_tokens - 1


def pad_block_table(block_ids, block_size: int, max_model_len: int, pad: int = BLOCK_TABLE_PAD):
    """RUNNER:746-752 / RUNNER:812-817 -- right-pad to max_model_len // block_size."""
    max_blocks_per_seq = max_model_len // block_size
    padded = [pad] * max_blocks_per_seq
    # the reference uses slice assignment, which GROWS the list when
    # len(block_ids) > max_blocks_per_seq; keep that behaviour.
    padded[: len(block_ids)] = list(block_ids)
    return padded


def prefill_index(prompt_len: int, block_ids, num_computed_tokens: int, block_size: int, max_model_len: int):
    """RUNNER:704-763 -- one new request (prefill, batch 1) with prefix caching.

    Returns the per-request rows exactly as the reference appends them to its
    IntermediateInputData: position_ids, full_context_len, padded block_table,
    computed_context_len and the slot_mapping slice that starts at the first
    uncached token.
    """
    block_table = list(block_ids)
    padded = pad_block_table(block_table, block_size, max_model_len)
    slot_mapping = [
        (block_table[i // block_size] * block_size + i % block_size) if i < prompt_len else SLOT_MAPPING_PAD
        for i in range(max_model_len)
    ]
    return {
        "position_ids": list(range(prompt_len)),
        "full_context_len": prompt_len,
        "computed_context_len": num_computed_tokens,
        "block_table": padded,
        "slot_mapping": slot_mapping[num_computed_tokens:],
    }


def decode_index(
    prompt_len: int,
    num_output_tokens: int,
    block_ids,
    block_size: int,
    max_model_len: int,
    tkg_kernel_pad: bool = False,
    num_speculative_tokens: int | None = None,
):
    """RUNNER:765-832 -- one cached (decode) request with prefix caching."""
    block_table = list(block_ids)
    pad = -1 if tkg_kernel_pad else BLOCK_TABLE_PAD  # RUNNER:805-811
    padded = pad_block_table(block_table, block_size, max_model_len, pad)
    position = last_token_position(prompt_len, num_output_tokens)
    slot = block_table[position // block_size] * block_size + position % block_size
    slots = [slot]
    if num_speculative_tokens is not None:  # RUNNER:825-830
        for i in range(1, num_speculative_tokens):
            slots.append(slots[0] + i)
    return {
        "position_ids": [position],
        "full_context_len": position + 1,
        "computed_context_len": position,
        "block_table": padded,
        "slot_mapping": slots,
    }


def _pad_rows(rows, pad, max_len):
    out = np.full((len(rows), max_len), pad, dtype=np.int64)
    for i, r in enumerate(rows):
        assert len(r) <= max_len
        out[i, : len(r)] = r
    return out


def finalize_inputs(rows: list[dict], input_tokens: list[list[int]], is_prefill: bool, max_model_len: int):
    """RUNNER:848-936 -- lists -> int64 arrays with the reference's pads/shapes."""
    full = [r["full_context_len"] for r in rows]
    if is_prefill:
        max_seq_len = max(full)
        assert max_seq_len > 0
        toks = _pad_rows(input_tokens, 0, max_seq_len)
        pos = _pad_rows([r["position_ids"] for r in rows], 0, max_seq_len)
        slots = _pad_rows([r["slot_mapping"] for r in rows], SLOT_MAPPING_PAD, max_model_len)
    else:
        toks = _pad_rows(input_tokens, 0, 1)
        pos = _pad_rows([r["position_ids"] for r in rows], 0, 1)
        slots = np.asarray([r["slot_mapping"] for r in rows], dtype=np.int64)
    return {
        "input_tokens": toks,
        "position_ids": pos,
        "slot_mapping": slots,
        "block_tables": np.asarray([r["block_table"] for r in rows], dtype=np.int64),
        "full_context_lens": np.asarray(full, dtype=np.int64).reshape(-1, 1),
        "computed_context_lens": np.asarray([r["computed_context_len"] for r in rows], dtype=np.int64).reshape(-1, 1),
    }


# ---------------------------------------------------------------- block counts
def null_block_adjust(num_gpu_blocks_override, already_adjusted: bool):
    """PLAT:149-159 -- +1 for vLLM's null block, once per CacheConfig."""
    if num_gpu_blocks_override is not None and not already_adjusted:
        return num_gpu_blocks_override + 1, True
    return num_gpu_blocks_override, already_adjusted


def default_num_blocks(max_model_len: int, block_size: int, max_num_seqs: int, num_gpu_blocks_override=None) -> int:
    """LOADER:741-745 -- note the floor inside the ceil (kept on purpose)."""
    n = ceil(max_model_len // block_size) * max_num_seqs
    if num_gpu_blocks_override is not None:
        n = num_gpu_blocks_override
    return n


def handle_pa_num_blocks(num_gpu_blocks_override, pa_num_blocks, explicitly_set: bool):
    """LOADER:796-831 -- returns the final pa_num_blocks or raises ValueError."""
    if num_gpu_blocks_override is not None:
        original_user_override = num_gpu_blocks_override - 1
        if explicitly_set:
            if pa_num_blocks == original_user_override:
                return num_gpu_blocks_override
            raise ValueError("pa_num_blocks must match --num-gpu-blocks-override intent")
        return pa_num_blocks
    if explicitly_set:
        raise ValueError("pa_num_blocks needs --num-gpu-blocks-override")
    return pa_num_blocks


def validate_block_count(
    is_block_kv_layout: bool, max_model_len: int, block_size: int, max_num_seqs: int, num_gpu_blocks_override, pa_num_blocks
):
    """LOADER:845-859 -- AssertionError when the pool is too small."""
    if not is_block_kv_layout:
        return
    min_blocks_required = ceil(max_model_len / block_size) * max_num_seqs
    if num_gpu_blocks_override is not None:
        effective = num_gpu_blocks_override - 1
    else:
        effective = pa_num_blocks
    assert effective >= min_blocks_required, (min_blocks_required, effective)


# ------------------------------------------------------------------ outputs
def strip_pad_tokens(sampled_rows):
    """RUNNER:304-306 -- drop -1 pads, keep 0 (a valid token id)."""
    return [[x for x in row if x != -1] for row in sampled_rows]


def sampling_params_rows(params, max_num_seqs: int, max_topk: int):
    """RUNNER:1106-1140 -- [max_num_seqs, 3] rows of (top_k, top_p, temperature).

    ``params`` is a list of (top_k, top_p, temperature) per live request in
    ``self.requests`` order; greedy (temperature 0) becomes (1, top_p, 1.0).
    """
    max_topk = min(max_topk, MAX_SAMPLING_TOP_K)
    top_k = [1] * max_num_seqs
    top_p = [1.0] * max_num_seqs
    temp = [1.0] * max_num_seqs
    for i, (k, p, t) in enumerate(params):
        top_k[i] = k if (k > 0 and k < max_topk) else max_topk
        top_p[i] = p
        temp[i] = t
        if t == 0.0:
            top_k[i] = 1
            temp[i] = 1.0
    return np.asarray([top_k, top_p, temp], dtype=np.float32).T.copy()


def remask_fused_spec_output(accepted, next_pos, positions):
    """LOADER:308-334 -- accepted tokens [B][T] arrive 0-padded from fused speculation; entries past the number of tokens
    generated this step (next_pos - position, clamped to [0, T]) become -1 so the runner strips them."""
    out = []
    for row, nxt, pos in zip(accepted, next_pos, positions):
        t = len(row)
        cnt = min(max(nxt - pos, 0), t)
        out.append(list(row[:cnt]) + [-1] * (t - cnt))
    return out
