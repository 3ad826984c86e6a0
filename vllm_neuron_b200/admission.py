"""Request admission: which sampling parameters the B200 device sampler can honour.

Import-light on purpose (no torch, no CUDA): the platform's front-end hook runs it in the API process for every request,
the model runner runs it again as a second line of defence.

The device sampler implements what the reference's on-device sampler takes -- (top_k, top_p, temperature) rows with
top_k <= 256 (RUNNER:1106-1140) -- plus seeds, min_tokens / stop handling in the scheduler, and log-probabilities.
Everything else vLLM's CPU Sampler would apply on the reference's CPU path (RUNNER:1142-1239) has no device implementation
here.  Such a request is refused at the FRONT END (B200Platform.validate_request -> only the offending client sees an
error); the reference's device path silently drops the same parameters (RUNNER:1106-1140 builds the [B,3] rows from
top_k / top_p / temperature alone).
"""
from __future__ import annotations

MAX_SAMPLING_TOP_K = 256  # RUNNER:74 (_MAX_NEURON_SAMPLING_TOP_K)

SUPPORTED = "temperature, top_k <= 256, top_p, seed, min_tokens, stop conditions, logprobs <= 256"


def unsupported_sampling_params(sp) -> list[str]:
    """Names (with values) of the parameters of `sp` that the device sampler cannot apply; [] when the request is fine."""
    if sp is None:
        return []
    bad = []
    for name, neutral in (("repetition_penalty", 1.0), ("presence_penalty", 0.0), ("frequency_penalty", 0.0), ("min_p", 0.0)):
        v = getattr(sp, name, neutral)
        if v is not None and float(v) != neutral:
            bad.append(f"{name}={v}")
    for name in ("logit_bias", "bad_words", "allowed_token_ids", "logits_processors", "structured_outputs", "guided_decoding"):
        v = getattr(sp, name, None)
        if v:
            bad.append(name)
    if getattr(sp, "prompt_logprobs", None) is not None:
        bad.append("prompt_logprobs")
    lp = getattr(sp, "logprobs", None)
    if lp is not None and not (0 <= int(lp) <= MAX_SAMPLING_TOP_K):
        bad.append(f"logprobs={lp} (0..{MAX_SAMPLING_TOP_K})")
    return bad


def check_sampling_params(sp, req_id="?") -> None:
    """Raises ValueError for a request the device sampler cannot serve.  Call it where an exception reaches only that
    request's client (the front end), never inside a worker step."""
    bad = unsupported_sampling_params(sp)
    if bad:
        raise ValueError(f"request {req_id}: sampling parameters not supported by the B200 device sampler: {', '.join(bad)} "
                         f"(supported: {SUPPORTED})")
