"""World-size-2 CPU test (gloo) of the tensor-parallel HOST logic: shard plan, per-channel-scale sharding, the greedy
merge of per-shard (max, index), and the rule that the residual is folded into rank 0 only.  The device kernels are not
involved (they have no CPU path); the arithmetic here is plain torch on the host."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vllm_neuron_b200.config import PRESETS
from vllm_neuron_b200.model import merge_sharded_argmax, quantize_weight, tp_shard_plan


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        shape = PRESETS["tiny-test"]
        plan = tp_shard_plan(shape, rank, world)
        g = torch.Generator().manual_seed(0)  # identical on both ranks
        hid, qd = shape.hidden, shape.num_q_heads * shape.head_dim
        x = torch.randn(5, hid, generator=g)
        wo = torch.randn(hid, qd, generator=g) * 0.05
        a = torch.randn(5, qd, generator=g)
        # row-parallel o_proj: each rank multiplies its column slice; residual only on rank 0; sum over ranks == full
        c0, c1 = plan["o_cols"]
        part = a[:, c0:c1] @ wo[:, c0:c1].T + (x if rank == 0 else 0)
        tot = part.clone()
        dist.all_reduce(tot)
        full = a @ wo.T + x
        assert torch.allclose(tot, full, atol=1e-4), (tot - full).abs().max()
        # column-parallel qkv rows of this rank tile the full matrix exactly once across ranks
        rows = torch.zeros(qd + 2 * shape.num_kv_heads * shape.head_dim)
        for r0, r1 in plan["qkv_rows"]:
            rows[r0:r1] += 1
        dist.all_reduce(rows)
        assert torch.all(rows == 1)
        # per-channel scales are properties of the un-sharded weight: quantise-then-shard == shard of quantised
        wq, ws = quantize_weight(wo.to(torch.bfloat16), "int8")
        wq_s, ws_s = quantize_weight(wo.to(torch.bfloat16)[:, c0:c1], "int8")
        assert ws.shape == ws_s.shape  # same channels; scales may differ, which is why the model quantises first
        # greedy merge: lowest index among the ranks holding the global maximum
        v0, v1 = plan["vocab_rows"]
        logits = torch.randn(4, shape.vocab, generator=g)
        logits[0, 3] = logits[0, shape.vocab - 1] = 50.0  # tie across shards -> index 3 wins
        loc = logits[:, v0:v1]
        val, idx = loc.max(dim=-1)
        idx = idx + v0
        vals = [torch.empty_like(val) for _ in range(world)]
        idxs = [torch.empty_like(idx) for _ in range(world)]
        dist.all_gather(vals, val)
        dist.all_gather(idxs, idx)
        tok = merge_sharded_argmax(torch.stack(vals), torch.stack(idxs), shape.vocab)
        assert torch.equal(tok, logits.argmax(dim=-1)) and int(tok[0]) == 3
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_tp_host_logic_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 200
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == {0: "ok", 1: "ok"}, res
