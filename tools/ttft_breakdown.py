import sys, torch
sys.path.insert(0, ".")
from vllm_neuron_b200 import ops
from vllm_neuron_b200.config import PRESETS
from vllm_neuron_b200.model import B200CausalLM
shape = PRESETS["llama-3.1-8b"]
n = 2048
mml = n + 32; mb = -(-mml // 32)
model = B200CausalLM(shape, None, quant="fp8", num_blocks=mb + 1, max_num_seqs=1, max_model_len=mml, device="cuda", seed=0)
model.block_table[0, :mb] = torch.arange(1, mb + 1, dtype=torch.int32, device="cuda")
toks = torch.randint(0, shape.vocab, (n,)).cuda()
model.prefill(0, toks, n, 0); torch.cuda.synchronize()
with ops.kernel_timer() as kt:
    model.prefill(0, toks, n, 0)
tot = 0
for name, (cnt, ms) in sorted(kt.summary().items(), key=lambda kv: -kv[1][0] * kv[1][1]):
    print(f"{name:45s} {cnt:4d} x {ms*1e3:9.1f} us = {cnt*ms:8.3f} ms"); tot += cnt * ms
print("sum", tot)
