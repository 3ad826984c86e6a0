"""Shared helpers for the parity tests (test infrastructure)."""
import torch


def logical_to_device_pool(pool_logical):
    """[nb, bs, Hkv, D] (oracle, token-major) -> [nb, Hkv, bs, D] (device layout)."""
    return pool_logical.permute(0, 2, 1, 3).contiguous()


def device_to_logical_pool(pool_dev):
    return pool_dev.permute(0, 2, 1, 3).contiguous()


def make_block_tables(b, ctx_lens, block_size, num_blocks, max_blocks, gen):
    """Random non-contiguous block ids from [1, num_blocks) (block 0 = null block), 0-padded rows."""
    need = [-(-int(c) // block_size) for c in ctx_lens]
    perm = (torch.randperm(num_blocks - 1, generator=gen) + 1).tolist()
    assert sum(need) <= len(perm)
    bt = torch.zeros((b, max_blocks), dtype=torch.int32)
    off = 0
    for i, n in enumerate(need):
        bt[i, :n] = torch.tensor(perm[off : off + n], dtype=torch.int32)
        off += n
    return bt


def rel_err(got, want):
    got, want = got.float(), want.float()
    return ((got - want).norm() / want.norm().clamp_min(1e-12)).item()


def max_rel_err(got, want, floor):
    got, want = got.float(), want.float()
    return ((got - want).abs() / want.abs().clamp_min(floor)).max().item()


# ---------------------------------------------------------------- duck-typed vLLM config / scheduler output
from types import SimpleNamespace  # noqa: E402


def hf_config_from_shape(shape):
    return SimpleNamespace(
        hidden_size=shape.hidden, num_hidden_layers=shape.layers, num_attention_heads=shape.num_q_heads,
        num_key_value_heads=shape.num_kv_heads, head_dim=shape.head_dim, intermediate_size=shape.intermediate,
        vocab_size=shape.vocab, rope_theta=shape.rope_theta, rope_scaling=shape.rope_scaling,
        rms_norm_eps=shape.rms_eps, attention_bias=shape.qkv_bias, architectures=["LlamaForCausalLM"],
        _name_or_path=shape.name,
    )


def make_vllm_config(shape, *, max_num_seqs, max_model_len, block_size=32, num_gpu_blocks_override=None,
                     prefix_caching=True, tp=1, override=None, enforce_eager=False):
    """The attributes of VllmConfig the runner reads (the reference's unit tests use Mock configs the same way,
    test/unit/conftest.py:104-145)."""
    return SimpleNamespace(
        model_config=SimpleNamespace(hf_config=hf_config_from_shape(shape), max_model_len=max_model_len, seed=0,
                                     enforce_eager=enforce_eager, dtype=torch.bfloat16),
        cache_config=SimpleNamespace(block_size=block_size, enable_prefix_caching=prefix_caching,
                                     num_gpu_blocks_override=num_gpu_blocks_override),
        parallel_config=SimpleNamespace(tensor_parallel_size=tp, world_size=tp),
        scheduler_config=SimpleNamespace(max_num_seqs=max_num_seqs, max_model_len=max_model_len,
                                         max_num_batched_tokens=131072),
        load_config=SimpleNamespace(load_format="dummy"),
        additional_config={"override_neuron_config": override or {}},
    )


def sched_output(new_reqs=(), cached=(), finished=()):
    """new_reqs: (req_id, prompt_ids, block_ids, num_computed); cached: (req_id, new_block_ids or None)."""
    num_sched = {}
    news = []
    for rid, prompt, blocks, ncomp in new_reqs:
        news.append(SimpleNamespace(req_id=rid, prompt_token_ids=list(prompt), block_ids=(list(blocks),),
                                    num_computed_tokens=ncomp, sampling_params=None, mm_features=[], lora_request=None))
        num_sched[rid] = len(prompt) - ncomp
    for rid, _ in cached:
        num_sched[rid] = 1
    c = SimpleNamespace(
        req_ids=[r for r, _ in cached], resumed_req_ids=set(),
        new_block_ids=[None if nb is None else (list(nb),) for _, nb in cached],
        num_computed_tokens=[0] * len(cached), new_token_ids=[], num_output_tokens=[0] * len(cached),
    )
    return SimpleNamespace(
        scheduled_new_reqs=news, scheduled_cached_reqs=c, num_scheduled_tokens=num_sched,
        total_num_scheduled_tokens=sum(num_sched.values()), finished_req_ids=set(finished),
        scheduled_spec_decode_tokens={}, free_encoder_mm_hashes=[],
    )


def hf_state_dict(full_weights, shape, *, lm_head=True):
    """The B200 weight dictionary (model.random_init_full_weights layout) under HF Llama/Qwen2 parameter names:
    fused qkv / gate-up rows split back into q/k/v and gate/up tensors."""
    qd, kd, I = shape.num_q_heads * shape.head_dim, shape.num_kv_heads * shape.head_dim, shape.intermediate
    sd = {"model.embed_tokens.weight": full_weights["embed"], "model.norm.weight": full_weights["final_norm"]}
    if lm_head:
        sd["lm_head.weight"] = full_weights["lm_head"]
    for li in range(shape.layers):
        lw, p = full_weights["layers"][li], f"model.layers.{li}."
        q, k, v = lw["wqkv"].split([qd, kd, kd])
        g, u = lw["wgu"].split([I, I])
        sd.update({p + "self_attn.q_proj.weight": q, p + "self_attn.k_proj.weight": k, p + "self_attn.v_proj.weight": v,
                   p + "self_attn.o_proj.weight": lw["wo"], p + "mlp.gate_proj.weight": g, p + "mlp.up_proj.weight": u,
                   p + "mlp.down_proj.weight": lw["wdown"], p + "input_layernorm.weight": lw["input_norm"],
                   p + "post_attention_layernorm.weight": lw["post_norm"]})
        if "bqkv" in lw:
            bq, bk, bv = lw["bqkv"].split([qd, kd, kd])
            sd.update({p + "self_attn.q_proj.bias": bq, p + "self_attn.k_proj.bias": bk, p + "self_attn.v_proj.bias": bv})
    return sd


def export_hf_checkpoint(full_weights, shape, out_dir, *, shards=2, lm_head=True):
    """Write the B200 weight dictionary as an HF Llama/Qwen2 safetensors checkpoint: separate q/k/v and gate/up tensors
    under the HF names, split over `shards` files + index json."""
    import json
    import os

    from safetensors.torch import save_file

    sd = hf_state_dict(full_weights, shape, lm_head=lm_head)
    os.makedirs(out_dir, exist_ok=True)
    names = sorted(sd)
    weight_map = {}
    for s in range(shards):
        fn = f"model-{s + 1:05d}-of-{shards:05d}.safetensors"
        part = {n: sd[n].contiguous().clone() for n in names[s::shards]}
        save_file(part, os.path.join(out_dir, fn))
        weight_map.update({n: fn for n in part})
    if shards > 1:
        with open(os.path.join(out_dir, "model.safetensors.index.json"), "w") as f:
            json.dump({"metadata": {}, "weight_map": weight_map}, f)
    return out_dir
