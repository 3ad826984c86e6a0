#!/usr/bin/env python
"""Generate tests/golden/index_golden.json by RUNNING THE REFERENCE'S OWN CODE.

Runs only in the authoring container (needs /root/reference, which does not exist
on the GPU box).  The reference's runner/loader modules do not import against the
installed vLLM 0.22 / without NxDI, so -- exactly like the reference's own unit
tests do (test/unit/conftest.py:15-51) -- the NxDI package tree is injected into
sys.modules as MagicMock, and the three vLLM 0.11 symbols that moved
(make_tensor_with_pad, MultiModalKwargs, set_random_seed) are provided as shims.
No reference source is copied: the functions are imported from where they lie
and called on synthetic inputs; only their integer outputs are stored.

    python tests/golden/gen_from_reference.py
"""
from __future__ import annotations

import json
import os
import random
import sys
import types
from types import SimpleNamespace
from unittest.mock import MagicMock

import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "index_golden.json")


def _install_shims():
    for name in [
        "neuronx_distributed_inference",
        "neuronx_distributed_inference.models",
        "neuronx_distributed_inference.models.config",
        "neuronx_distributed_inference.models.mllama",
        "neuronx_distributed_inference.models.mllama.utils",
        "neuronx_distributed_inference.models.llama4",
        "neuronx_distributed_inference.models.llama4.utils",
        "neuronx_distributed_inference.models.llama4.utils.input_processor",
        "neuronx_distributed_inference.modules",
        "neuronx_distributed_inference.modules.generation",
        "neuronx_distributed_inference.modules.generation.sampling",
        "neuronx_distributed_inference.modules.lora_serving",
        "neuronx_distributed_inference.modules.padding",
        "neuronx_distributed_inference.utils",
        "neuronx_distributed_inference.utils.constants",
        "neuronx_distributed_inference.utils.hf_adapter",
        "neuronx_distributed",
        "neuronx_distributed.quantization",
        "neuronx_distributed.quantization.quantization_config",
        "torch_neuronx",
    ]:
        sys.modules.setdefault(name, MagicMock())

    import vllm.utils as vu

    if not hasattr(vu, "make_tensor_with_pad"):

        def make_tensor_with_pad(x, pad, dtype, *, max_len=None, device=None, pin_memory=False):
            # vLLM 0.11 vllm/utils/__init__.py semantics: right-pad each row to max_len.
            max_len = max_len if max_len is not None else max(map(len, x), default=0)
            rows = []
            for r in x:
                r = list(r)
                assert len(r) <= max_len
                rows.append(r + [pad] * (max_len - len(r)))
            return torch.tensor(rows, dtype=dtype, device=device).reshape(len(rows), max_len)

        vu.make_tensor_with_pad = make_tensor_with_pad
    import vllm.multimodal as vmm

    if not hasattr(vmm, "MultiModalKwargs"):
        vmm.MultiModalKwargs = MagicMock()
    import vllm.model_executor as vme

    if not hasattr(vme, "set_random_seed"):
        vme.set_random_seed = lambda seed: None
    sys.path.insert(0, REF)


def _mk_runner(RUNNER, block_size, max_model_len, max_num_seqs, tkg=False, spec=None):
    r = object.__new__(RUNNER.NeuronxDistributedModelRunner)
    r.cache_config = SimpleNamespace(block_size=block_size)
    r.scheduler_config = SimpleNamespace(max_model_len=max_model_len, max_num_seqs=max_num_seqs)
    r.speculative_config = None if spec is None else SimpleNamespace(num_speculative_tokens=spec)
    r.lora_config = None
    r.is_prefix_caching = True
    r.is_chunked_prefill = False
    r.requests = {}
    r.vllm_req_to_neuron_seq_id_mapping = {}
    r.free_seq_ids = set(range(max_num_seqs))
    r.device = torch.device("cpu")
    r.model = SimpleNamespace(
        neuron_config=SimpleNamespace(
            attn_tkg_nki_kernel_enabled=tkg, attn_block_tkg_nki_kernel_enabled=False, on_device_sampling_config=None
        )
    )
    r.get_nxd_sampling_params = lambda input_ids: None
    return r


def _tolist(t):
    return None if t is None else t.tolist()


def gen_index_cases(RUNNER):
    rng = random.Random(20261104)
    cases = []
    cfgs = [(32, 2048, 4), (32, 256, 4), (16, 96, 2), (32, 64, 3), (8, 64, 5)]
    for block_size, max_model_len, max_num_seqs in cfgs:
        num_blocks = max(64, (max_model_len // block_size) * max_num_seqs + 1)
        # ---- prefill, batch 1 (SCHED:116), with and without a prefix-cache hit
        for trial in range(4):
            runner = _mk_runner(RUNNER, block_size, max_model_len, max_num_seqs)
            prompt_len = [1, block_size, max_model_len - 1, rng.randint(2, max_model_len - 2)][trial]
            nblk = -(-prompt_len // block_size)
            block_ids = rng.sample(range(1, num_blocks), nblk)
            ncomp = 0 if trial % 2 == 0 else (rng.randint(0, (prompt_len - 1) // block_size) * block_size)
            prompt = [rng.randint(0, 31999) for _ in range(prompt_len)]
            req = SimpleNamespace(
                req_id=f"r{trial}",
                prompt_token_ids=prompt,
                block_ids=(list(block_ids),),
                num_computed_tokens=ncomp,
                mm_features=None,
                lora_request=None,
            )
            data = RUNNER.IntermediateInputData()
            runner._process_new_request_for_continuous_batching(req, data)
            mi = runner._finalize_continuous_batching_inputs(data, True)
            cases.append(
                {
                    "kind": "prefill",
                    "block_size": block_size,
                    "max_model_len": max_model_len,
                    "max_num_seqs": max_num_seqs,
                    "prompt_token_ids": prompt,
                    "block_ids": block_ids,
                    "num_computed_tokens": ncomp,
                    "out": {
                        "input_tokens": _tolist(mi.input_tokens),
                        "position_ids": _tolist(mi.position_ids),
                        "slot_mapping": _tolist(mi.slot_mapping),
                        "block_tables": _tolist(mi.block_tables),
                        "full_context_lens": _tolist(mi.full_context_lens),
                        "computed_context_lens": _tolist(mi.computed_context_lens),
                        "input_block_ids": _tolist(mi.input_block_ids),
                    },
                }
            )
        # ---- decode batch (all cached requests)
        for tkg, spec in [(False, None), (True, None), (False, 3)]:
            runner = _mk_runner(RUNNER, block_size, max_model_len, max_num_seqs, tkg=tkg, spec=spec)
            reqs = []
            for i in range(max_num_seqs):
                prompt_len = rng.randint(1, max_model_len - 8)
                nout = rng.randint(1, min(6, max_model_len - prompt_len - 1))
                ntok = prompt_len + nout
                nblk = -(-ntok // block_size)
                block_ids = rng.sample(range(1, num_blocks), nblk)
                rid = f"d{i}"
                outs = [rng.randint(0, 31999) for _ in range(nout)]
                runner.requests[rid] = SimpleNamespace(
                    prompt_token_ids=[0] * prompt_len, output_token_ids=outs, block_ids=(list(block_ids),)
                )
                runner.vllm_req_to_neuron_seq_id_mapping[rid] = runner.free_seq_ids.pop()
                reqs.append({"req_id": rid, "prompt_len": prompt_len, "output_token_ids": outs, "block_ids": block_ids})
            cached = SimpleNamespace(req_ids=[q["req_id"] for q in reqs])
            data = RUNNER.IntermediateInputData()
            for i in range(len(reqs)):
                runner._process_cached_request_for_continuous_batching(cached, i, data)
            mi = runner._finalize_continuous_batching_inputs(data, False)
            cases.append(
                {
                    "kind": "decode",
                    "block_size": block_size,
                    "max_model_len": max_model_len,
                    "max_num_seqs": max_num_seqs,
                    "tkg": tkg,
                    "spec": spec,
                    "reqs": reqs,
                    "out": {
                        "input_tokens": _tolist(mi.input_tokens),
                        "position_ids": _tolist(mi.position_ids),
                        "slot_mapping": _tolist(mi.slot_mapping),
                        "block_tables": _tolist(mi.block_tables),
                        "full_context_lens": _tolist(mi.full_context_lens),
                        "computed_context_lens": _tolist(mi.computed_context_lens),
                        "input_block_ids": _tolist(mi.input_block_ids),
                    },
                }
            )
    return cases


def gen_config_cases(LOADER):
    """LOADER:725-867 on a grid of (max_model_len, block_size, max_num_seqs, override, pa_num_blocks)."""
    out = []
    grid = []
    for mml, bs, mns in [(2048, 32, 4), (1000, 16, 10), (2048, 32, 64), (128, 32, 2), (4096, 32, 128)]:
        need = -(-mml // bs) * mns
        for override in [None, need, need + 1, need + 1000, need - 1, 4096]:
            for pa in [None, "same", "plus1", 7]:
                grid.append((mml, bs, mns, override, pa))
    for mml, bs, mns, user_override, pa in grid:
        # platform step (PLAT:149-159): +1 once
        adj = None if user_override is None else user_override + 1
        cache = SimpleNamespace(block_size=bs, num_gpu_blocks_override=adj, enable_prefix_caching=True)
        sched = SimpleNamespace(max_model_len=mml, max_num_seqs=mns, chunked_prefill_enabled=False, max_num_batched_tokens=131072)
        par = SimpleNamespace(tensor_parallel_size=1)
        mc = SimpleNamespace(dtype=torch.bfloat16)
        rec = {"max_model_len": mml, "block_size": bs, "max_num_seqs": mns, "user_override": user_override, "pa": pa}
        try:
            cfg = LOADER._get_default_neuron_config(mc, cache, par, sched, None, None)
            rec["default_pa_num_blocks"] = cfg["pa_num_blocks"]
            rec["pa_block_size"] = cfg["pa_block_size"]
            rec["is_block_kv_layout"] = bool(cfg["is_block_kv_layout"])
            rec["is_prefix_caching"] = bool(cfg["is_prefix_caching"])
            rec["batch_size"] = cfg["batch_size"]
            rec["tp_degree"] = cfg["tp_degree"]
            rec["seq_len"] = cfg["seq_len"]
            rec["max_context_length"] = cfg["max_context_length"]
            ov = {}
            if pa is not None:
                base = user_override if user_override is not None else 100
                ov["pa_num_blocks"] = {"same": base, "plus1": base + 1, 7: 7}[pa]
            rec["override"] = dict(ov)
            cfg = LOADER._get_neuron_config_after_override(cfg, dict(ov))
            cfg = LOADER._handle_pa_num_blocks(cache, cfg, ov)
            LOADER._validate_neuron_config(cache, sched, cfg)
            rec["final_pa_num_blocks"] = cfg["pa_num_blocks"]
            rec["error"] = None
        except (ValueError, AssertionError) as e:
            rec["error"] = type(e).__name__
        out.append(rec)
    return out


def gen_platform_cases():
    """PLAT:131-210 on SimpleNamespace configs (the reference's own tests use Mocks the same way)."""
    from vllm_neuron.platform import NeuronPlatform

    NeuronPlatform._config_overrides_applied = True  # do not monkey-patch ModelConfig in this process
    out = []
    for override in [None, 0, 100, 4096]:
        for prefix in [True, False]:
            for world in [1, 8]:
                cache = SimpleNamespace(num_gpu_blocks_override=override, enable_prefix_caching=prefix, block_size=32)
                par = SimpleNamespace(worker_cls="auto", world_size=world, distributed_executor_backend=None)
                sched = SimpleNamespace(scheduler_cls=None, chunked_prefill_enabled=True, max_num_batched_tokens=8192, max_num_seqs=0)
                cfg = SimpleNamespace(
                    model_config=SimpleNamespace(max_model_len=2048), cache_config=cache, parallel_config=par, scheduler_config=sched
                )
                NeuronPlatform.check_and_update_config(cfg)
                NeuronPlatform.check_and_update_config(cfg)  # idempotence of the +1
                out.append(
                    {
                        "override_in": override,
                        "prefix": prefix,
                        "world": world,
                        "override_out": cache.num_gpu_blocks_override,
                        "block_size_out": cache.block_size,
                        "max_num_seqs_out": sched.max_num_seqs,
                        "max_num_batched_tokens_out": sched.max_num_batched_tokens,
                        "chunked_prefill_out": sched.chunked_prefill_enabled,
                        "worker_cls_is_set": par.worker_cls != "auto",
                        "scheduler_cls_is_set": sched.scheduler_cls is not None,
                        "backend_out": par.distributed_executor_backend,
                    }
                )
    return out


def gen_sampling_cases(RUNNER):
    """RUNNER:1106-1140: capture the top_k/top_p/temperature lists handed to NxDI."""
    out = []
    rng = random.Random(7)
    for mns in [4, 8]:
        for vocab, global_topk in [(32000, 256), (100, 64), (128256, 1000)]:
            runner = _mk_runner(RUNNER, 32, 2048, mns)
            del runner.get_nxd_sampling_params
            runner.is_chunked_prefill = False
            runner.model_config = SimpleNamespace(get_vocab_size=lambda v=vocab: v)
            runner.model.neuron_config.on_device_sampling_config = SimpleNamespace(global_topk=global_topk)
            params = []
            for i in range(rng.randint(1, mns)):
                k = rng.choice([-1, 0, 1, 5, 50, 255, 256, 300, 2000])
                p = rng.choice([1.0, 0.9, 0.5])
                t = rng.choice([0.0, 0.7, 1.0, 1.3])
                params.append([k, p, t])
                runner.requests[f"s{i}"] = SimpleNamespace(sampling_params=SimpleNamespace(top_k=k, top_p=p, temperature=t))
            captured = {}

            def fake_prepare(batch_size, top_k, top_p, temperature):
                captured.update(batch_size=batch_size, top_k=list(top_k), top_p=list(top_p), temperature=list(temperature))
                return torch.tensor([top_k, top_p, temperature], dtype=torch.float32).T

            RUNNER.prepare_sampling_params = fake_prepare
            runner.get_nxd_sampling_params(torch.zeros(len(params), 1, dtype=torch.long))
            out.append({"max_num_seqs": mns, "vocab": vocab, "global_topk": global_topk, "params": params, **captured})
    return out


def gen_output_cases(RUNNER):
    """RUNNER:293-359: -1 stripping of sampled ids (no speculation)."""
    out = []
    for rows in [[[5], [0], [-1]], [[-1], [-1]], [[7, -1], [0, 3]], [[31999]]]:
        runner = _mk_runner(RUNNER, 32, 2048, 4)
        n = len(rows)
        runner.max_model_len = 2048
        runner.input_batch = SimpleNamespace(
            num_tokens_no_spec=[10] * n,
            num_tokens=[10] * n,
            token_ids_cpu=torch.zeros(n, 2048, dtype=torch.int64).numpy(),
            req_ids=[f"o{i}" for i in range(n)],
            req_id_to_index={f"o{i}": i for i in range(n)},
        )
        for i in range(n):
            runner.requests[f"o{i}"] = SimpleNamespace(output_token_ids=[])
        so = SimpleNamespace(sampled_token_ids=torch.tensor(rows), logprobs_tensors=None)
        mro = runner._generate_model_runner_output(so)
        out.append(
            {
                "rows": rows,
                "sampled_token_ids": [list(map(int, r)) for r in mro.sampled_token_ids],
                "num_tokens_after": list(map(int, runner.input_batch.num_tokens)),
                "output_token_ids": [list(map(int, runner.requests[f"o{i}"].output_token_ids)) for i in range(n)],
            }
        )
    return out


def gen_remask_cases(LOADER):
    """LOADER:308-334 (_remask_fused_spec_output): 0-padded accepted tokens -> -1 past the tokens generated this step."""
    out = []
    rng = random.Random(11)
    for b, t in [(1, 1), (3, 4), (4, 8), (2, 3)]:
        accepted = [[rng.randint(0, 31999) for _ in range(t)] for _ in range(b)]
        positions = [rng.randint(0, 2000) for _ in range(b)]
        counts = [rng.choice([-2, 0, 1, t - 1, t, t + 3]) for _ in range(b)]
        next_pos = [p + c for p, c in zip(positions, counts)]
        fused = [torch.tensor(accepted), torch.zeros(b, 1), torch.tensor(next_pos).unsqueeze(-1)]
        inputs = {"position_ids": torch.tensor([[0, p] for p in positions])}
        masked = LOADER.NeuronCausalLM._remask_fused_spec_output(None, fused, inputs)
        out.append({"accepted": accepted, "positions": positions, "next_pos": next_pos, "masked": _tolist(masked)})
    return out


def gen_scheduler_cases():
    """SCHED:59-166 run on plain objects: the stop rule on a grid of request states, can_schedule on a grid of queue sizes,
    and schedule()'s queue movement around the (stubbed) base-scheduler call."""
    from collections import deque

    import vllm_neuron.core.scheduler as SCHED
    from vllm.v1.request import RequestStatus

    stop = []
    rng = random.Random(5)
    for _ in range(160):
        n_out = rng.randint(1, 12)
        prompt = rng.randint(1, 20)
        eos = 2
        last = rng.choice([2, 2, 7, 9, 11])
        req = SimpleNamespace(
            num_tokens=prompt + n_out, num_output_tokens=n_out, max_tokens=rng.choice([n_out - 1, n_out, n_out + 1, 64]),
            pooling_params=None, eos_token_id=eos, output_token_ids=[5] * (n_out - 1) + [last], status=None, stop_reason=None,
            sampling_params=SimpleNamespace(min_tokens=rng.choice([0, 0, n_out, n_out + 2]), ignore_eos=rng.choice([False, False, True]),
                                            stop_token_ids=rng.choice([None, [], [7], [9, 11]]), eos_token_id=eos))
        mml = rng.choice([prompt + n_out - 1, prompt + n_out, prompt + n_out + 1, 4096])
        rec = {"num_tokens": req.num_tokens, "num_output_tokens": n_out, "max_tokens": req.max_tokens, "eos_token_id": eos,
               "last_token": last, "min_tokens": req.sampling_params.min_tokens, "ignore_eos": req.sampling_params.ignore_eos,
               "stop_token_ids": req.sampling_params.stop_token_ids, "max_model_len": mml}
        rec["stopped"] = bool(SCHED.check_stop_with_min_tokens(req, mml))
        rec["status"] = None if req.status is None else RequestStatus(req.status).name
        rec["stop_reason"] = req.stop_reason
        stop.append(rec)

    can = []
    for max_run in (1, 2, 4):
        for n_run in range(0, max_run + 1):
            for n_wait in range(0, 3):
                fake = SimpleNamespace(running=[0] * n_run, waiting=[0] * n_wait, max_num_running_reqs=max_run,
                                       scheduler_config=SimpleNamespace(max_model_len=2048))
                can.append({"max_running": max_run, "running": n_run, "waiting": n_wait,
                            "can": bool(SCHED.ContinuousBatchingNeuronScheduler.can_schedule(fake, None))})

    sched = []
    seen = {}

    def base_schedule(self):  # stands in for vLLM's Scheduler.schedule: records what it is shown
        seen["waiting"], seen["running"] = list(self.waiting), list(self.running)
        return "OUT"

    orig = SCHED.Scheduler.schedule
    SCHED.Scheduler.schedule = base_schedule
    try:
        for max_run in (1, 2, 4, 8):
            for n_run in range(0, max_run + 1):
                for n_wait in (0, 1, 2, 5):
                    obj = object.__new__(SCHED.ContinuousBatchingNeuronScheduler)
                    obj.waiting = deque(f"w{i}" for i in range(n_wait))
                    obj.running = [f"r{i}" for i in range(n_run)]
                    obj.holdback_queue = deque()
                    obj.max_num_running_reqs = max_run
                    obj.scheduler_config = SimpleNamespace(max_model_len=2048)
                    out = obj.schedule()
                    assert out == "OUT"
                    sched.append({"max_running": max_run, "running": n_run, "waiting": n_wait,
                                  "base_saw_waiting": seen["waiting"], "base_saw_running": seen["running"],
                                  "waiting_after": list(obj.waiting), "running_after": list(obj.running)})
    finally:
        SCHED.Scheduler.schedule = orig
    return {"stop": stop, "can_schedule": can, "schedule": sched}


def main():
    _install_shims()
    import vllm_neuron.worker.neuronx_distributed_model_loader as LOADER
    import vllm_neuron.worker.neuronx_distributed_model_runner as RUNNER

    golden = {
        "_generated_by": "tests/golden/gen_from_reference.py (runs the reference's own functions)",
        "index": gen_index_cases(RUNNER),
        "config": gen_config_cases(LOADER),
        "platform": gen_platform_cases(),
        "sampling": gen_sampling_cases(RUNNER),
        "output": gen_output_cases(RUNNER),
        "remask": gen_remask_cases(LOADER),
        "scheduler": gen_scheduler_cases(),
    }
    with open(OUT, "w") as f:
        json.dump(golden, f, separators=(",", ":"))
    print(OUT, os.path.getsize(OUT), "bytes;", {k: len(v) for k, v in golden.items() if isinstance(v, list)})


if __name__ == "__main__":
    main()
