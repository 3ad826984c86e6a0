"""Debug: run one prefill through ops step by step, comparing every intermediate with the oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import kv_ref, quant_ref, attn_ref, index_ref
from vllm_neuron_b200 import ops
from vllm_neuron_b200.config import PRESETS
from vllm_neuron_b200.model import B200CausalLM, random_init_full_weights

quant = sys.argv[1] if len(sys.argv) > 1 else "fp8"
T = int(sys.argv[2]) if len(sys.argv) > 2 else 70
shape = PRESETS["tiny-test"]
bs, mml, nb = 32, 256, 12
w = random_init_full_weights(shape, seed=3, std=0.05)
m = B200CausalLM(shape, w, quant=quant, num_blocks=nb, max_num_seqs=2, max_model_len=mml)
g = torch.Generator().manual_seed(0)
toks = torch.randint(0, shape.vocab, (T,), generator=g)
blocks = [3, 7, 5]
m.block_table[0, :3] = torch.tensor(blocks, dtype=torch.int32, device="cuda")
row = index_ref.prefill_index(T, blocks, 0, bs, mml)
slots = torch.tensor(row["slot_mapping"][:T]); pos = torch.arange(T)

def cmp(name, got, want):
    got = got.float().cpu(); want = want.float()
    err = (got - want).abs().max().item(); ref = want.abs().max().item()
    nz = ((got - want).abs() > 1e-2 * ref).float().mean().item()
    print(f"{name:14s} max|err| {err:.3e}  max|ref| {ref:.3e}  frac>1% {nz:.4f}")

L = m.layers[0]; lw = w["layers"][0]
x_d = ops.embedding(toks.cuda(), m.embed); x_r = w["embed"][toks]
cmp("embed", x_d, x_r)
qk = None if quant == "bf16" else quant
def lin(lin_d, w_full, xd, xr, residual_d=None, residual_r=None, bias=None):
    if qk is None:
        yd = ops.gemm(xd, lin_d.w, None, None, lin_d.bias, residual_d, workspace=m.gemm_ws)
        yr = quant_ref.gemm_ref(xr, w_full, bias, residual_r)
    else:
        if xd.dtype == torch.bfloat16:
            xq, xs = ops.quantize_per_token(xd, qk)
        else:
            xq, xs = xd
        yd = ops.gemm(xq, lin_d.w, xs, lin_d.scale, lin_d.bias, residual_d, workspace=m.gemm_ws)
        wq, ws = quant_ref.quantize_weight_per_channel(w_full, qk)
        cmp("  wq", lin_d.w.float(), wq.float()); cmp("  ws", lin_d.scale, ws)
        xrq, xrs = quant_ref.quantize_act_per_token(xr, qk)
        cmp("  xq", xq.float(), xrq.float()); cmp("  xs", xs, xrs)
        yr = quant_ref.qgemm_ref(xrq, xrs, wq, ws, bias, residual_r)
        yself = quant_ref.qgemm_ref(xq.cpu(), xs.cpu(), wq, ws, bias, residual_r)
        cmp("  gemm(self)", yd, yself)
    return yd, yr
h_d = ops.rmsnorm(x_d, L["input_norm"], shape.rms_eps); h_r = kv_ref.rmsnorm_ref(x_r, lw["input_norm"], shape.rms_eps)
cmp("norm1", h_d, h_r)
qkv_d, qkv_r = lin(L["qkv"], lw["wqkv"], h_d, h_r)
cmp("qkv", qkv_d, qkv_r)
qd, kd = m.hq * m.d, m.hkv * m.d
cos, sin = kv_ref.rope_table(m.d, mml, shape.rope_theta, None)
q_r = kv_ref.rope_ref(qkv_r[:, :qd].reshape(T, m.hq, m.d), pos, cos, sin)
k_r = kv_ref.rope_ref(qkv_r[:, qd:qd+kd].reshape(T, m.hkv, m.d), pos, cos, sin)
v_r = qkv_r[:, qd+kd:].reshape(T, m.hkv, m.d)
kp = torch.zeros(nb, bs, m.hkv, m.d, dtype=torch.bfloat16); vp = torch.zeros_like(kp)
kv_ref.kv_write_ref(k_r, v_r, kp, vp, slots)
ops.rope_kv_write(qkv_d, pos.cuda(), m.cos, m.sin, m.kv[0][0], m.kv[0][1], slots.cuda(), m.hq)
cmp("q roped", qkv_d[:, :qd], q_r.reshape(T, qd))
cmp("k pool", m.kv[0][0].permute(0, 2, 1, 3), kp)
a_d = ops.paged_attn_prefill(qkv_d[:, :qd].unflatten(1, (m.hq, m.d)), m.kv[0][0], m.kv[0][1], m.block_table[0], 0, m.scale)
a_r = attn_ref.paged_attention_prefill_ref(q_r, kp, vp, torch.tensor(row["block_table"]), 0, m.scale)
cmp("attn", a_d, a_r)
x2_d, x2_r = lin(L["o"], lw["wo"], a_d.reshape(T, qd), a_r.reshape(T, qd), x_d, x_r)
cmp("o+res", x2_d, x2_r)
h2_d = ops.rmsnorm(x2_d, L["post_norm"], shape.rms_eps); h2_r = kv_ref.rmsnorm_ref(x2_r, lw["post_norm"], shape.rms_eps)
cmp("norm2", h2_d, h2_r)
gu_d, gu_r = lin(L["gu"], lw["wgu"], h2_d, h2_r)
cmp("gate_up", gu_d, gu_r)
act_d = ops.silu_mul(gu_d); act_r = kv_ref.silu_mul_ref(gu_r)
cmp("act", act_d, act_r)
x3_d, x3_r = lin(L["down"], lw["wdown"], act_d, act_r, x2_d, x2_r)
cmp("down+res", x3_d, x3_r)
