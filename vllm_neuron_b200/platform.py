"""B200Platform: the out-of-tree vLLM platform, in the shape of the reference's NeuronPlatform.

Reference: vllm_neuron/platform.py
  class attributes                         :25-31
  check_and_update_config                  :131-210  (null block +1 :149-159, worker_cls :162-164,
                                                      scheduler :176-187, budget/defaults :189-201, block_size rule :203-210)
  is_pin_memory_available/use_all_gather/supports_v1  :212-223

Deliberate B200 differences (DESIGN.md "boundary"):
  * device_type "cuda", dist_backend "nccl", CUDA_VISIBLE_DEVICES; pinned memory IS available.
  * the executor is NOT forced to "uni" (PLAT:166-167): tensor parallelism is one worker process per GPU.
  * ModelConfig verifiers are not monkey-patched (PLAT:42-113): the head-divisibility check is wanted here.
  * one attention backend, no selection logic: get_attn_backend_cls always names the block-KV paged backend.
"""
from __future__ import annotations

import logging
import os
from typing import TYPE_CHECKING

from vllm.platforms import Platform, PlatformEnum

if TYPE_CHECKING:
    from vllm.config import VllmConfig

logger = logging.getLogger(__name__)

WORKER_CLS = "vllm_neuron_b200.worker.b200_worker.B200Worker"
SCHEDULER_CLS = "vllm_neuron_b200.core.scheduler.ContinuousBatchingB200Scheduler"
ATTN_BACKEND_CLS = "vllm_neuron_b200.attention.B200PagedAttentionBackend"
_NULL_BLOCK_MARK = "_b200_null_block_adjusted"
DEFAULT_MAX_NUM_BATCHED_TOKENS = 131072  # PLAT:190
DEFAULT_MAX_NUM_SEQS = 32  # PLAT:198


class B200Platform(Platform):
    _enum = PlatformEnum.OOT
    device_name: str = "cuda"
    device_type: str = "cuda"
    dispatch_key: str = "CUDA"
    ray_device_key: str = "GPU"
    dist_backend: str = "nccl"
    device_control_env_var: str = "CUDA_VISIBLE_DEVICES"
    # reference names kept ("neuron_quant", "fbgemm_fp8") + the native ones
    supported_quantization: list[str] = ["neuron_quant", "fbgemm_fp8", "fp8", "int8", "b200_w8a8"]

    @classmethod
    def get_device_name(cls, device_id: int = 0) -> str:
        return "b200"

    @classmethod
    def is_async_output_supported(cls, enforce_eager: bool | None) -> bool:
        return False  # PLAT:119-121

    @classmethod
    def pre_register_and_update(cls, parser=None) -> None:
        return None

    @classmethod
    def check_and_update_config(cls, vllm_config: "VllmConfig") -> None:
        model_config = vllm_config.model_config
        if model_config is None:  # vLLM validates empty VllmConfig objects too (PLAT:135-144)
            return
        disable_scheduler_override = bool(int(
            os.getenv("DISABLE_B200_CUSTOM_SCHEDULER", os.getenv("DISABLE_NEURON_CUSTOM_SCHEDULER", "0"))))

        # +1 for vLLM's lazily allocated null block, once per CacheConfig instance (PLAT:149-159)
        cache_config = vllm_config.cache_config
        if cache_config and cache_config.num_gpu_blocks_override is not None \
                and _NULL_BLOCK_MARK not in cache_config.__dict__:
            logger.info("Adding 1 to num_gpu_blocks_override (%d -> %d) to account for null block allocation",
                        cache_config.num_gpu_blocks_override, cache_config.num_gpu_blocks_override + 1)
            cache_config.num_gpu_blocks_override += 1
            setattr(cache_config, _NULL_BLOCK_MARK, True)

        parallel_config = vllm_config.parallel_config
        if parallel_config.worker_cls == "auto":
            parallel_config.worker_cls = WORKER_CLS
        # (no "uni" override: one process per GPU over NCCL)

        sched_cfg = vllm_config.scheduler_config
        if disable_scheduler_override:
            logger.warning("The vLLM V1 native scheduler will be used with chunked prefill enabled; the B200 decode "
                           "kernels assume prefill batch 1 and no mixed steps, so this path is not the fast one.")
            assert cache_config.block_size is not None, "When vLLM V1 native scheduler is enabled, block_size must be set."
        else:
            sched_cfg.scheduler_cls = SCHEDULER_CLS
            # vLLM 0.11 names the switch chunked_prefill_enabled (PLAT:181), 0.2x enable_chunked_prefill
            for name in ("chunked_prefill_enabled", "enable_chunked_prefill"):
                if hasattr(sched_cfg, name):
                    try:
                        setattr(sched_cfg, name, False)
                    except Exception:  # a read-only property in some versions
                        pass
            sched_cfg.max_num_batched_tokens = DEFAULT_MAX_NUM_BATCHED_TOKENS
            if getattr(sched_cfg, "async_scheduling", False):
                # vLLM >= 0.2x turns async scheduling on by default; the prefill-or-decode policy needs each step's
                # tokens and finishes before the next schedule() (the reference is synchronous too, PLAT:119-121)
                logger.info("Disabling async scheduling: the continuous-batching scheduler is synchronous")
                sched_cfg.async_scheduling = False
            if not sched_cfg.max_num_seqs:
                sched_cfg.max_num_seqs = DEFAULT_MAX_NUM_SEQS
            if not cache_config.enable_prefix_caching:
                # the reference falls back to one block per sequence (PLAT:203-207); the B200 kernels are block-KV only
                # (block_size 32), so the rule is kept only when the user left block_size unset
                if cache_config.block_size is None:
                    cache_config.block_size = model_config.max_model_len
            else:
                assert cache_config.block_size is not None, "When prefix caching is enabled, block_size must be set."

    @classmethod
    def validate_request(cls, processed_inputs=None, params=None, *args, **kwargs) -> None:
        """Front-end admission (vLLM's input processor calls this once per request, in the API process): refuse sampling
        parameters the device sampler cannot apply, so that only the offending client gets an error.  The reference has
        no such hook: its device path drops those parameters silently (RUNNER:1106-1140), its CPU path applies them
        (RUNNER:1142-1239); there is no CPU path here."""
        if params is None or not hasattr(params, "temperature"):  # PoolingParams and the like: not this platform's call
            return
        from .admission import check_sampling_params

        check_sampling_params(params, getattr(processed_inputs, "request_id", None) or "?")

    @classmethod
    def get_attn_backend_cls(cls, selected_backend=None, *args, **kwargs) -> str:
        return ATTN_BACKEND_CLS  # one backend; no dispatch

    @classmethod
    def get_device_communicator_cls(cls) -> str:
        # torch.distributed (NCCL) on the TP device group; the hot-path exchange is this plugin's own NVLink kernel
        # (tp_comm.py), so vLLM's CudaCommunicator (custom all-reduce, asserts a builtin CUDA platform) is not wanted
        return "vllm.distributed.device_communicators.base_device_communicator.DeviceCommunicatorBase"

    @classmethod
    def is_pin_memory_available(cls) -> bool:
        return True

    @classmethod
    def use_all_gather(cls) -> bool:
        return True  # PLAT:217-219

    @classmethod
    def supports_v1(cls, model_config=None) -> bool:
        return True  # PLAT:221-223

    @classmethod
    def get_device_capability(cls, device_id: int = 0):
        from vllm.platforms.interface import DeviceCapability

        return DeviceCapability(major=10, minor=0)

    @classmethod
    def get_device_total_memory(cls, device_id: int = 0) -> int:
        import torch

        return torch.cuda.get_device_properties(device_id).total_memory

    @classmethod
    def set_device(cls, device) -> None:
        import torch

        torch.cuda.set_device(device)
