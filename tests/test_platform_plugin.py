"""CPU tests of the plugin boundary: register(), B200Platform.check_and_update_config, worker surface.
Modelled on the reference's test/unit/test_init.py:9-41 and test/unit/test_platform.py:156-248, 673-762."""
import warnings
from types import SimpleNamespace

import pytest

import vllm_neuron_b200 as pkg


def test_register_without_device_warns_and_returns_none(monkeypatch):
    monkeypatch.setattr(pkg, "_sm100_present", lambda: False)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        assert pkg.register() is None
    assert any(issubclass(x.category, UserWarning) and "B200" in str(x.message) for x in w)


def test_register_with_device_returns_platform_qualname(monkeypatch):
    monkeypatch.setattr(pkg, "_sm100_present", lambda: True)
    assert pkg.register() == "vllm_neuron_b200.platform.B200Platform"


def test_entry_point_declared():
    import os

    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "setup.py")).read()
    assert '"vllm.platform_plugins"' in src and "vllm_neuron_b200:register" in src


def _cfg(**kw):
    cache = SimpleNamespace(num_gpu_blocks_override=kw.get("override"), block_size=kw.get("block_size", 32),
                            enable_prefix_caching=kw.get("prefix", True))
    return SimpleNamespace(
        model_config=SimpleNamespace(max_model_len=2048),
        cache_config=cache,
        parallel_config=SimpleNamespace(worker_cls=kw.get("worker_cls", "auto"), world_size=kw.get("world", 1),
                                        distributed_executor_backend=kw.get("backend", None)),
        scheduler_config=SimpleNamespace(scheduler_cls=None, chunked_prefill_enabled=True, max_num_batched_tokens=8192,
                                         max_num_seqs=kw.get("max_num_seqs", 0)),
    )


@pytest.fixture(scope="module")
def platform():
    from vllm_neuron_b200.platform import B200Platform

    return B200Platform


def test_class_attributes(platform):
    from vllm.platforms import PlatformEnum

    assert platform._enum == PlatformEnum.OOT
    assert platform.device_type == "cuda" and platform.dist_backend == "nccl"
    assert platform.device_control_env_var == "CUDA_VISIBLE_DEVICES"
    for q in ("neuron_quant", "fbgemm_fp8", "fp8", "int8"):
        assert q in platform.supported_quantization
    assert platform.supports_v1(None) and platform.use_all_gather()
    assert platform.is_async_output_supported(None) is False
    assert platform.get_attn_backend_cls(None) == "vllm_neuron_b200.attention.B200PagedAttentionBackend"


def test_empty_config_is_ignored(platform):
    cfg = _cfg()
    cfg.model_config = None
    platform.check_and_update_config(cfg)  # must not raise (PLAT:135-144)
    assert cfg.parallel_config.worker_cls == "auto"


def test_null_block_added_once(platform):
    cfg = _cfg(override=4096)
    platform.check_and_update_config(cfg)
    assert cfg.cache_config.num_gpu_blocks_override == 4097  # README.md:37,45 of the reference: 4096 -> 4097
    platform.check_and_update_config(cfg)  # same CacheConfig instance: idempotent (test_platform.py:673-762)
    assert cfg.cache_config.num_gpu_blocks_override == 4097
    cfg2 = _cfg(override=None)
    platform.check_and_update_config(cfg2)
    assert cfg2.cache_config.num_gpu_blocks_override is None


def test_worker_and_scheduler_selection(platform, monkeypatch):
    monkeypatch.delenv("DISABLE_NEURON_CUSTOM_SCHEDULER", raising=False)
    monkeypatch.delenv("DISABLE_B200_CUSTOM_SCHEDULER", raising=False)
    cfg = _cfg(world=8, backend="mp")
    platform.check_and_update_config(cfg)
    assert cfg.parallel_config.worker_cls == "vllm_neuron_b200.worker.b200_worker.B200Worker"
    assert cfg.parallel_config.distributed_executor_backend == "mp"  # NOT forced to "uni": one process per GPU
    assert cfg.scheduler_config.scheduler_cls == "vllm_neuron_b200.core.scheduler.ContinuousBatchingB200Scheduler"
    assert cfg.scheduler_config.chunked_prefill_enabled is False
    assert cfg.scheduler_config.max_num_batched_tokens == 131072
    assert cfg.scheduler_config.max_num_seqs == 32
    keep = _cfg(worker_cls="my.Worker", max_num_seqs=4)
    platform.check_and_update_config(keep)
    assert keep.parallel_config.worker_cls == "my.Worker" and keep.scheduler_config.max_num_seqs == 4


def test_scheduler_override_disabled_by_env(platform, monkeypatch):
    monkeypatch.setenv("DISABLE_NEURON_CUSTOM_SCHEDULER", "1")
    cfg = _cfg()
    platform.check_and_update_config(cfg)
    assert cfg.scheduler_config.scheduler_cls is None and cfg.scheduler_config.max_num_batched_tokens == 8192


def test_block_size_rules(platform, monkeypatch):
    monkeypatch.delenv("DISABLE_NEURON_CUSTOM_SCHEDULER", raising=False)
    cfg = _cfg(prefix=False, block_size=None)
    platform.check_and_update_config(cfg)
    assert cfg.cache_config.block_size == 2048  # PLAT:203-207
    with pytest.raises(AssertionError):
        platform.check_and_update_config(_cfg(prefix=True, block_size=None))  # PLAT:208-210


def test_worker_surface():
    from vllm_neuron_b200.worker.b200_worker import B200Worker

    for m in ("init_device", "load_model", "get_kv_cache_spec", "determine_available_memory", "initialize_from_config",
              "initialize_cache", "compile_or_warm_up_model", "execute_model", "check_health", "get_supported_tasks",
              "take_draft_token_ids", "add_lora", "remove_lora", "pin_lora", "list_loras", "profile", "sample_tokens",
              "get_cache_block_size_bytes", "shutdown"):
        assert callable(getattr(B200Worker, m)), m


def test_async_scheduling_is_switched_off(platform, monkeypatch):
    """vLLM 0.22 enables async scheduling by default (two batches in flight, tokens taken from sample_tokens); the
    prefill-or-decode policy is synchronous like the reference's (PLAT:119-121), so the platform turns it off."""
    monkeypatch.delenv("DISABLE_NEURON_CUSTOM_SCHEDULER", raising=False)
    monkeypatch.delenv("DISABLE_B200_CUSTOM_SCHEDULER", raising=False)
    cfg = _cfg()
    cfg.scheduler_config.async_scheduling = True
    platform.check_and_update_config(cfg)
    assert cfg.scheduler_config.async_scheduling is False


def test_worker_hands_the_step_output_to_sample_tokens():
    """Batch-queue engine loops read the ModelRunnerOutput from sample_tokens(), the plain loop from execute_model()."""
    from vllm_neuron_b200.worker.b200_worker import B200Worker

    w = B200Worker.__new__(B200Worker)
    w.is_driver_worker = True
    w.model_runner = SimpleNamespace(execute_model=lambda so: ("out", so))
    assert w.execute_model("s1") == ("out", "s1")
    assert w.sample_tokens(None) == ("out", "s1") and w.sample_tokens(None) is None
    w.is_driver_worker = False
    assert w.execute_model("s2") is None and w.sample_tokens(None) is None  # WORKER:65-69


def test_device_communicator_class_resolves(platform):
    """TP>1 under vLLM's multi-process executor builds the TP GroupCoordinator with this class; CudaCommunicator asserts a
    builtin CUDA platform (custom all-reduce), the base communicator is plain torch.distributed on the device group."""
    import importlib

    mod, _, name = platform.get_device_communicator_cls().rpartition(".")
    cls = getattr(importlib.import_module(mod), name)
    assert hasattr(cls, "all_reduce") and hasattr(cls, "all_gather")


def test_validate_request_refuses_what_the_device_sampler_lacks(platform):
    """vLLM's input processor calls Platform.validate_request(processed_inputs, params) per request in the front end:
    the place where a ValueError reaches one client instead of the engine core."""
    from vllm import SamplingParams
    from vllm.pooling_params import PoolingParams

    platform.validate_request({}, SamplingParams(temperature=0.8, top_k=40, top_p=0.9, seed=1, logprobs=5, min_tokens=2))
    platform.validate_request({}, SamplingParams())
    platform.validate_request({}, PoolingParams())  # not a sampling request: nothing to say
    for kw, word in (({"repetition_penalty": 1.1}, "repetition_penalty=1.1"), ({"presence_penalty": 0.5}, "presence_penalty"),
                     ({"frequency_penalty": 0.5}, "frequency_penalty"), ({"min_p": 0.1}, "min_p"),
                     ({"logit_bias": {3: 1.0}}, "logit_bias"), ({"allowed_token_ids": [1, 2]}, "allowed_token_ids"),
                     ({"prompt_logprobs": 1}, "prompt_logprobs"), ({"logprobs": 300}, "logprobs=300")):
        with pytest.raises(ValueError, match=word):
            platform.validate_request({}, SamplingParams(**kw))


def test_input_processor_reaches_the_platform_hook():
    """The installed vLLM really calls the hook from its input processor (otherwise the refusal above is dead code)."""
    import inspect

    from vllm.v1.engine import input_processor

    assert "current_platform.validate_request(" in inspect.getsource(input_processor)
