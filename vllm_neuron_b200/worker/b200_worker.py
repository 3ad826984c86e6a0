"""B200Worker: vLLM worker for one GPU, in the shape of the reference's NeuronWorker.

Reference: vllm_neuron/worker/neuron_worker.py
  ctor                              :25-43      execute_model (driver-only return)   :65-69
  init_device / distributed init    :45-49,106-121   (gloo, world size 1, ceremonial)
  determine_available_memory        :51-63      get_kv_cache_spec / initialize_from_config :95-100
  LoRA stubs / tasks / draft tokens :123-146
Here the process group is real: one worker per GPU, NCCL, world size = tensor_parallel_size.
"""
from __future__ import annotations

import logging
import os

import torch

logger = logging.getLogger(__name__)

try:  # vLLM present (GPU box, real engine)
    from vllm.v1.worker.worker_base import WorkerBase
except Exception:  # pragma: no cover - unit tests without vLLM
    class WorkerBase:  # type: ignore
        def __init__(self, vllm_config, local_rank, rank, distributed_init_method, is_driver_worker=False):
            self.vllm_config = vllm_config
            self.model_config = vllm_config.model_config
            self.cache_config = vllm_config.cache_config
            self.parallel_config = vllm_config.parallel_config
            self.local_rank, self.rank = local_rank, rank
            self.distributed_init_method = distributed_init_method
            self.is_driver_worker = is_driver_worker


class B200Worker(WorkerBase):
    def __init__(self, vllm_config, local_rank: int, rank: int, distributed_init_method: str,
                 is_driver_worker: bool = False) -> None:
        super().__init__(vllm_config=vllm_config, local_rank=local_rank, rank=rank,
                         distributed_init_method=distributed_init_method, is_driver_worker=is_driver_worker)
        self.device = torch.device("cuda", local_rank)
        self.model_runner = None
        self._tp_group = None

    # ------------------------------------------------------------------ lifecycle
    def init_device(self) -> None:
        from .. import ops

        torch.cuda.set_device(self.device)
        ops.device_check(self.device.index)  # raises if this is not an sm_100 part / the CUDA library is missing
        tp = self.parallel_config.tensor_parallel_size
        self._init_distributed(tp)
        seed = getattr(self.model_config, "seed", None)
        if seed is not None:
            torch.manual_seed(seed)
        from .b200_model_runner import B200ModelRunner

        self.model_runner = B200ModelRunner(self.vllm_config, self.device, tp_rank=self.rank % tp, tp_size=tp,
                                            tp_group=self._tp_group)

    def _init_distributed(self, tp: int) -> None:
        """WORKER:106-121 initialises a ceremonial gloo group of size 1; here: NCCL, world size = TP."""
        try:
            from vllm.distributed import ensure_model_parallel_initialized, init_distributed_environment

            init_distributed_environment(world_size=self.parallel_config.world_size, rank=self.rank,
                                         local_rank=self.local_rank,
                                         distributed_init_method=self.distributed_init_method, backend="nccl")
            ensure_model_parallel_initialized(tp, 1)
            if tp > 1:
                from vllm.distributed import get_tp_group

                self._tp_group = get_tp_group().device_group
        except ImportError:
            if tp > 1 and not torch.distributed.is_initialized():
                torch.distributed.init_process_group("nccl", init_method=self.distributed_init_method,
                                                     world_size=tp, rank=self.rank, device_id=self.device)
            self._tp_group = torch.distributed.group.WORLD if tp > 1 else None

    def load_model(self, *, load_dummy_weights: bool = False) -> None:
        self.model_runner.load_model(load_dummy_weights=load_dummy_weights)

    def determine_available_memory(self) -> int:
        """Bytes the KV pools may take: free HBM x gpu_memory_utilization minus the (not yet materialised) weights and the
        reserve of the shared CUDA-graph pool.
        (The reference asks the Neuron runtime or assumes 20 GB, WORKER:51-63.)"""
        free, total = torch.cuda.mem_get_info(self.device)
        util = getattr(self.cache_config, "gpu_memory_utilization", 0.9)
        budget = (int(total * util) - (total - free) - self.model_runner.weight_bytes_estimate()
                  - self.model_runner.graph_pool_reserve_bytes())
        return max(budget, 0)

    def get_kv_cache_spec(self):
        return self.model_runner.get_kv_cache_spec()

    def get_cache_block_size_bytes(self) -> int:
        shape = self.model_runner.model_shape()
        tp = self.parallel_config.tensor_parallel_size
        return 2 * self.cache_config.block_size * (shape.num_kv_heads // tp) * shape.head_dim * 2 * shape.layers

    def initialize_cache(self, num_gpu_blocks: int, num_cpu_blocks: int) -> None:
        self.cache_config.num_gpu_blocks = num_gpu_blocks  # WORKER:79-83
        self.cache_config.num_cpu_blocks = num_cpu_blocks

    def initialize_from_config(self, kv_cache_config) -> None:
        self.model_runner.initialize_kv_cache(kv_cache_config)

    def compile_or_warm_up_model(self):
        """Captures the decode CUDA graph for the full batch so the first request does not pay for it."""
        self.model_runner.warm_up()
        try:  # vLLM >= 0.2x: the executor takes max() over the workers' CompilationTimes
            from vllm.v1.worker.worker_base import CompilationTimes

            return CompilationTimes(language_model=0.0, encoder=0.0)
        except ImportError:
            return None

    def update_max_model_len(self, max_model_len: int) -> None:
        self.model_config.max_model_len = max_model_len
        if self.model_runner is not None:
            self.model_runner.max_model_len = max_model_len

    # ------------------------------------------------------------------ the step
    def execute_model(self, scheduler_output):
        output = self.model_runner.execute_model(scheduler_output)
        self._last_output = output if self.is_driver_worker else None
        return self._last_output  # WORKER:65-69

    def sample_tokens(self, grammar_output=None):
        """Sampling already happened on the device inside execute_model.  vLLM's batch-queue engine loop ignores
        execute_model's return value and takes the step's ModelRunnerOutput from here instead."""
        out, self._last_output = getattr(self, "_last_output", None), None
        return out

    def take_draft_token_ids(self):
        return None

    # ------------------------------------------------------------------ misc contract
    def check_health(self) -> None:
        return

    def profile(self, is_start: bool = True):
        if is_start:
            torch.cuda.profiler.start()
        else:
            torch.cuda.profiler.stop()

    def get_model(self):
        return self.model_runner.model

    def get_supported_tasks(self):
        return ("generate",)

    def add_lora(self, lora_request) -> bool:
        raise NotImplementedError("LoRA is outside the B200 hot path (SURVEY 8: out of scope)")

    def remove_lora(self, lora_id: int) -> bool:
        return False

    def pin_lora(self, lora_id: int) -> bool:
        return False

    def list_loras(self):
        return set()

    def shutdown(self) -> None:
        self.model_runner = None
        torch.cuda.empty_cache()
