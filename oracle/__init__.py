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
* Arithmetic side (``kv_ref``, ``attn_ref``, ``quant_ref``, ``llama_ref``):
  PARITY UNPINNED.  The reference delegates every floating-point op to
  neuronx-distributed-inference (NxDI, AWS Neuron SDK 2.26.1), which is neither
  vendored under /root/reference nor installable here, and no reference test pins
  a numeric attention/GEMM/KV result.  These files restate the published
  algorithms (block-table gather + softmax(QK^T/sqrt(D))V in fp32; symmetric
  per-channel quantisation; Llama decoder) from first principles.
"""
