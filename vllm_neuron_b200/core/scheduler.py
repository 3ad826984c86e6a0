"""Continuous-batching scheduler policy the B200 kernels rely on: a step is EITHER one prefill (batch 1) OR a
decode batch of every running request, never a mix.

Reference: vllm_neuron/core/scheduler.py
  ContinuousBatchingNeuronScheduler.schedule / can_schedule   :69-127
  _update_request_with_output + check_stop_with_min_tokens    :32-56, 130-166

The policy itself (`admit_count`, `stop_reason`) is plain Python over counts so it is unit-testable without an
engine; the vLLM subclass only moves requests between queues around the base scheduler's `schedule()`.
"""
from __future__ import annotations

from collections import deque

MAX_PROMPT_BATCH_SIZE = 1  # SCHED:116


def admit_count(num_running: int, num_pending: int, max_num_running_reqs: int,
                max_prompt_batch: int = MAX_PROMPT_BATCH_SIZE) -> int:
    """How many of the pending requests (FIFO) may enter this step's prefill (SCHED:82-90, 115-127):
    a request is admitted if nothing is running or waiting at all, or if the decode batch still has a free line
    AND the prefill batch (size 1) is not full."""
    admitted = 0
    for _ in range(num_pending):
        start_new_batch = num_running + admitted == 0
        has_line = num_running + admitted < max_num_running_reqs
        prefill_free = admitted < max_prompt_batch
        if start_new_batch or (has_line and prefill_free):
            admitted += 1
        else:
            break
    return admitted


def stop_reason(*, num_tokens: int, num_output_tokens: int, max_tokens: int, max_model_len: int, min_tokens: int,
                last_token_id: int, eos_token_id, ignore_eos: bool, stop_token_ids=(), pooling: bool = False,
                has_pooler_output: bool = False):
    """SCHED:130-166 as a pure function.  Returns None (keep going), "length", "stop" or ("stop", token_id).
    Length caps win over min_tokens; min_tokens suppresses EOS / stop tokens until reached."""
    if num_tokens >= max_model_len or num_output_tokens >= max_tokens:
        return "length"
    if pooling:
        return "stop" if has_pooler_output else None
    if min_tokens > 0 and num_output_tokens < min_tokens:
        return None
    if not ignore_eos and last_token_id == eos_token_id:
        return "stop"
    if last_token_id in (stop_token_ids or ()):
        return ("stop", last_token_id)
    return None


def _vllm_scheduler_base():
    from vllm.v1.core.sched.scheduler import Scheduler

    return Scheduler


_REQUEST_STATUS = None


def check_stop_with_min_tokens(request, max_model_len: int, pooler_output=None) -> bool:
    """SCHED:130-166 on a vLLM Request (same signature): `stop_reason`'s rule written out on the object's fields -- this
    runs once per request per step inside the engine loop, so no keyword packing and no per-call import."""
    global _REQUEST_STATUS
    if _REQUEST_STATUS is None:
        from vllm.v1.request import RequestStatus

        _REQUEST_STATUS = RequestStatus
    sp = request.sampling_params
    pooling = request.pooling_params
    if not pooling:
        assert sp is not None
    n_out = request.num_output_tokens
    if request.num_tokens >= max_model_len or n_out >= request.max_tokens:
        request.status = _REQUEST_STATUS.FINISHED_LENGTH_CAPPED
        return True
    if pooling:
        if pooler_output is not None:
            request.status = _REQUEST_STATUS.FINISHED_STOPPED
            return True
        return False
    min_tokens = sp.min_tokens
    if min_tokens > 0 and n_out < min_tokens:
        return False
    out = request.output_token_ids
    last = out[-1] if out else -1
    if not sp.ignore_eos:
        # vLLM 0.11 keeps the EOS id on the Request (SCHED:149), 0.2x on its SamplingParams (None when ignore_eos)
        eos = request.eos_token_id if hasattr(request, "eos_token_id") else getattr(sp, "eos_token_id", None)
        if last == eos:
            request.status = _REQUEST_STATUS.FINISHED_STOPPED
            return True
    stop_ids = sp.stop_token_ids
    if stop_ids and last in stop_ids:
        request.status = _REQUEST_STATUS.FINISHED_STOPPED
        request.stop_reason = last
        return True
    return False


def _build_scheduler_cls():
    Base = _vllm_scheduler_base()

    class ContinuousBatchingB200Scheduler(Base):
        def __init__(self, *args, **kwargs) -> None:
            super().__init__(*args, **kwargs)
            self.holdback_queue: deque = deque()

        def _update_request_with_output(self, request, new_token_ids):
            stopped = False
            for n, tok in enumerate(new_token_ids, 1):
                request.append_output_token_ids(tok)
                stopped = check_stop_with_min_tokens(request, self.max_model_len)
                if stopped:
                    del new_token_ids[n:]
                    break
            return new_token_ids, stopped

        def can_schedule(self, request) -> bool:
            return admit_count(len(self.running) + len(self.waiting), 1, self.max_num_running_reqs,
                               MAX_PROMPT_BATCH_SIZE - len(self.waiting)) == 1

        def schedule(self):
            pending = deque()
            while self.waiting:
                pending.append(self.waiting.pop_request() if hasattr(self.waiting, "pop_request") else self.waiting.popleft())
            n = admit_count(len(self.running), len(pending), self.max_num_running_reqs)
            for _ in range(n):
                self._push_waiting(pending.popleft())
            self.holdback_queue = pending
            hidden_running = []
            if len(self.waiting) > 0:  # prefill step: hide the decodes from the base scheduler
                hidden_running, self.running = self.running, []
            out = super().schedule()
            self.running = self.running + hidden_running
            while self.holdback_queue:
                self._push_waiting(self.holdback_queue.popleft())
            return out

        def _push_waiting(self, req):
            if hasattr(self.waiting, "add_request"):
                self.waiting.add_request(req)
            else:
                self.waiting.append(req)

    return ContinuousBatchingB200Scheduler


def __getattr__(name):
    if name == "ContinuousBatchingB200Scheduler":
        cls = _build_scheduler_cls()
        globals()[name] = cls
        return cls
    raise AttributeError(name)
