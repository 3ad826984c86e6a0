"""CPU oracle for the block-KV paged decode hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``vllm_neuron_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` do, and there only as the checker or
the timed CPU baseline -- never as the product path.

Parity status
-------------
* Integer / index / config side (``index_ref``): PINNED.  The functions restate
  the reference's Python (file:line cited per function) and are checked against
  golden vectors produced by importing the reference's own modules in the
  authoring container (``tests/golden/gen_from_reference.py``).
* Arithmetic side (``kv_ref``, ``attn_ref``, ``quant_ref``, ``llama_ref``): PINNED TO HF TRANSFORMERS (round 2).
  The reference delegates every floating-point op to neuronx-distributed-inference (NxDI, AWS Neuron SDK 2.26.1),
  which is neither vendored under /root/reference nor installable here, and no reference test pins a numeric
  attention/GEMM/KV result; the reference's own accuracy method is greedy generation with HF transformers on CPU
  (test/e2e/online/online_server_runner.py:95-146).  ``tests/golden/gen_from_hf.py`` runs exactly that -- HF's
  LlamaForCausalLM / Qwen2ForCausalLM, fp32, CPU -- on seeded checkpoints (tiny-test, tiny-qwen, the TinyLlama-1.1B
  shape) and commits logits + 64 greedy tokens as ``tests/golden/hf_golden.json``; ``tests/test_oracle_hf_pin.py``
  asserts the oracle against them: with fp32 activations it reproduces HF's logits to < 1e-5 of the logit spread and
  all 64 tokens through both its block-KV prefill and paged decode paths (algorithm pinned: RoPE convention, GQA
  mapping, norm placement, block-table gather, causal mask, SiLU gating); INT8 weight-only equals HF holding the
  de-quantised weights; the W8A8 modes equal HF + a per-token fake-quant hook on every nn.Linear until the first 8-bit
  rounding flip; the bf16 oracle the GPU tests use is the same code with bf16 rounding points (distance to HF measured
  and bounded).  What stays WITHOUT an external reference: the quantisation RULE (scale = amax / qmax, e4m3fn at 448;
  the reference only forwards config keys to NxDI, LOADER:885-898) and the device RNG of the sampler.
* Sampler (``sampling_ref``): the kept-token set and probabilities are pinned to vLLM's own top-k / top-p filter
  (what the reference's CPU sampling path runs, RUNNER:1142-1239), ``tests/test_oracle_sampling.py``.
"""
