"""Run ContinuousBatchingB200Scheduler inside the REAL vLLM scheduler/KV-manager on the CPU (no GPU, no model): requests go
in, a fake model runner answers every scheduled request with one token, and the sequence of SchedulerOutputs is printed
as JSON.  tests/test_scheduler_engine_cpu.py asserts the policy the kernels rely on (SCHED:69-127: a step is one prefill
or a decode batch, never a mix; a waiting request enters when a batch line is free) and that B200Platform's config
rewrite (scheduler class, null block +1, async scheduling off) really reaches a VllmConfig built by vLLM's EngineArgs.
The plugin is activated through its entry point with the device probe forced true (there is no device here)."""
from __future__ import annotations

import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
os.environ.setdefault("VLLM_PLUGINS", "b200")
os.environ.setdefault("VLLM_LOGGING_LEVEL", "ERROR")


def main() -> None:
    import vllm_engine_e2e as e2e

    e2e.ensure_entry_point()
    import vllm_neuron_b200

    vllm_neuron_b200._sm100_present = lambda: True
    import torch

    from vllm_neuron_b200.config import PRESETS

    shape = PRESETS["tiny-test"]
    d = tempfile.mkdtemp(prefix="b200_sched_")
    with open(os.path.join(d, "config.json"), "w") as f:
        json.dump({"architectures": ["LlamaForCausalLM"], "model_type": "llama", "hidden_size": shape.hidden,
                   "num_hidden_layers": shape.layers, "num_attention_heads": shape.num_q_heads,
                   "num_key_value_heads": shape.num_kv_heads, "head_dim": shape.head_dim,
                   "intermediate_size": shape.intermediate, "vocab_size": shape.vocab, "rope_theta": 10000.0,
                   "rms_norm_eps": 1e-5, "max_position_embeddings": 4096, "torch_dtype": "bfloat16",
                   "tie_word_embeddings": False, "hidden_act": "silu", "bos_token_id": 1, "eos_token_id": 2}, f)
    from vllm.engine.arg_utils import EngineArgs

    vc = EngineArgs(model=d, skip_tokenizer_init=True, dtype="bfloat16", load_format="dummy", max_model_len=256,
                    max_num_seqs=4, block_size=32, enable_prefix_caching=True,
                    num_gpu_blocks_override=48).create_engine_config()
    from vllm.platforms import current_platform
    from vllm.utils.hashing import sha256
    from vllm.v1.core.kv_cache_utils import get_request_block_hasher, init_none_hash
    from vllm.v1.kv_cache_interface import FullAttentionSpec, KVCacheConfig, KVCacheGroupSpec
    from vllm.v1.outputs import ModelRunnerOutput
    from vllm.v1.request import Request
    from vllm.v1.structured_output import StructuredOutputManager

    from vllm import SamplingParams
    from vllm_neuron_b200.core.scheduler import ContinuousBatchingB200Scheduler

    res = {"platform": type(current_platform).__name__, "scheduler_cls": str(vc.scheduler_config.scheduler_cls),
           "async_scheduling": bool(vc.scheduler_config.async_scheduling), "worker_cls": vc.parallel_config.worker_cls,
           "num_gpu_blocks_override": vc.cache_config.num_gpu_blocks_override,
           "chunked_prefill": bool(getattr(vc.scheduler_config, "enable_chunked_prefill", False))}
    vc.cache_config.num_gpu_blocks = vc.cache_config.num_gpu_blocks_override
    spec = FullAttentionSpec(block_size=32, num_kv_heads=shape.num_kv_heads, head_size=shape.head_dim, dtype=torch.bfloat16)
    kvc = KVCacheConfig(num_blocks=vc.cache_config.num_gpu_blocks, kv_cache_tensors=[],
                        kv_cache_groups=[KVCacheGroupSpec([f"layers.{i}.attn" for i in range(shape.layers)], spec)])
    sched = ContinuousBatchingB200Scheduler(vllm_config=vc, kv_cache_config=kvc,
                                            structured_output_manager=StructuredOutputManager(vc), block_size=32)
    init_none_hash(sha256)
    hasher = get_request_block_hasher(32, sha256)

    def req(name, tokens, max_tokens):
        return Request(request_id=name, prompt_token_ids=tokens, pooling_params=None, block_hasher=hasher,
                       sampling_params=SamplingParams(temperature=0, max_tokens=max_tokens, ignore_eos=True))

    base = [(7 * j) % 1000 + 3 for j in range(70)]
    # six requests for four batch lines; r1 stops early; r5 repeats r0's first 64 tokens (two full blocks)
    plan = [("r0", base, 6), ("r1", [5] * 33, 3), ("r2", [9] * 5, 6), ("r3", [8] * 40, 6), ("r4", [4] * 20, 6),
            ("r5", base[:64] + [17] * 9, 4)]
    for name, toks, mt in plan:
        sched.add_request(req(name, toks, mt))
    steps, produced = [], {name: 0 for name, _, _ in plan}
    for _ in range(200):
        if not sched.has_requests():
            break
        so = sched.schedule()
        new = [(r.req_id, r.num_computed_tokens, [list(b) for b in r.block_ids]) for r in so.scheduled_new_reqs]
        cached = list(so.scheduled_cached_reqs.req_ids)
        steps.append({"new": new, "cached": cached, "num_scheduled_tokens": dict(so.num_scheduled_tokens),
                      "finished": sorted(so.finished_req_ids)})
        ids = [n[0] for n in new] + cached
        for r in ids:
            produced[r] += 1
        out = ModelRunnerOutput(req_ids=ids, req_id_to_index={r: i for i, r in enumerate(ids)},
                                sampled_token_ids=[[11] for _ in ids], logprobs=None, prompt_logprobs_dict={}, pooler_output=[])
        sched.update_from_output(so, out)
    res.update(steps=steps, produced=produced, plan={n: mt for n, _, mt in plan}, max_num_seqs=4)
    print("RESULT " + json.dumps(res))


if __name__ == "__main__":
    main()
