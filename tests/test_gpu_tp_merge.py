"""GPU (one device): b200_tp_argmax_merge, the greedy merge over the vocabulary shards of a tensor-parallel group.

A single GPU plays one rank of a `world`-rank group: the other ranks' (max, index) pairs are ordinary device arrays, and
their barrier flags are pre-set to a far-future epoch, so the kernel's "wait until every peer has signalled" passes at
once.  What is checked is the merge rule itself -- the largest value wins, ties go to the lowest global index, exactly
model.merge_sharded_argmax (the TP=1 argmax rule) -- for ragged batch sizes, forced ties across shards, repeated replays
on one buffer, and rows whose maxima are NaN (the token must stay a valid index).  The real cross-GPU barrier is covered
by tests/test_gpu_tp.py and tools/tp_check.py --stress."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
FAR = 0x7FFFFFFF


@pytest.fixture(scope="module")
def ops():
    from vllm_neuron_b200 import ops as o

    o.device_check(0)
    return o


class _FakeGroup:
    def __init__(self, world, rank, rows):
        dev = "cuda"
        self.world, self.rank = world, rank
        self.vals = [torch.zeros(rows, dtype=torch.float32, device=dev) for _ in range(world)]
        self.idxs = [torch.zeros(rows, dtype=torch.int64, device=dev) for _ in range(world)]
        # flags[p] = rank p's flag words, one per peer; only OUR words are waited on: peers "signalled" long ago
        self.flags = [torch.zeros(8, dtype=torch.int32, device=dev) for _ in range(world)]
        self.flags[rank].fill_(FAR)
        self.epoch = torch.zeros(1, dtype=torch.int32, device=dev)
        arr = C.c_void_p * world
        self.pv = arr(*[t.data_ptr() for t in self.vals])
        self.pi = arr(*[t.data_ptr() for t in self.idxs])
        self.pf = arr(*[t.data_ptr() for t in self.flags])

    def load(self, vals, idxs):
        for p in range(self.world):
            self.vals[p][: vals.shape[1]].copy_(vals[p])
            self.idxs[p][: idxs.shape[1]].copy_(idxs[p])


@pytest.mark.parametrize("world,rank", [(1, 0), (2, 1), (4, 0), (8, 5)])
def test_merge_rule_ragged_batches_ties_and_replays(ops, world, rank):
    from vllm_neuron_b200.model import merge_sharded_argmax

    vocab, shard = 8 * 1000, 1000
    grp = _FakeGroup(world, rank, 64)
    g = torch.Generator().manual_seed(world * 10 + rank)
    step = 0
    for b in (1, 7, 33, 64, 7, 64):  # shrinking and growing again: stale pairs of earlier, larger batches stay behind
        vals = torch.randn(world, b, generator=g)
        idxs = torch.stack([torch.randint(0, shard, (b,), generator=g) + p * shard for p in range(world)])
        if world > 1:
            # forced ties: every other row has the same maximum on several shards (the lowest index must win) ...
            top = vals.max(dim=0).values
            for r in range(0, b, 2):
                k = 2 + (r // 2) % (world - 1)
                for p in torch.randperm(world, generator=g)[:k].tolist():
                    vals[p, r] = top[r]
            # ... and one row is tied on ALL shards at -inf (an all-masked row)
            vals[:, b - 1] = float("-inf")
        grp.load(vals, idxs)
        want = merge_sharded_argmax(vals, idxs, vocab)
        for _ in range(3):  # replays on the same buffers: the epoch advances, the answer does not
            got = ops.tp_argmax_merge(grp.pv, grp.pi, grp.pf, grp.epoch, rank, world, b)
            step += 1
            torch.cuda.synchronize()
            assert torch.equal(got.cpu(), want), (world, b)
        assert int(grp.epoch.item()) == step
        # our signal reached every "peer": their flag word for this rank carries the epoch
        assert all(int(grp.flags[p][rank].item()) == step for p in range(world) if p != rank)


def test_merge_of_nan_maxima_stays_a_valid_token(ops):
    world, b = 4, 5
    grp = _FakeGroup(world, 2, 8)
    nan = float("nan")
    vals = torch.tensor([[nan, nan, 1.0, nan, -1.0],
                         [nan, 2.0, nan, nan, nan],
                         [nan, nan, 3.0, nan, -1.0],
                         [nan, 2.0, nan, 0.5, nan]])
    idxs = torch.tensor([[10, 11, 12, 13, 14], [110, 111, 112, 113, 114], [210, 211, 212, 213, 214], [310, 311, 312, 313, 314]])
    grp.load(vals, idxs)
    got = ops.tp_argmax_merge(grp.pv, grp.pi, grp.pf, grp.epoch, 2, world, b).cpu().tolist()
    # all NaN -> rank 0's index (in range, never INT64_MAX); otherwise NaN loses to any number, ties to the lowest index
    assert got == [10, 111, 212, 313, 14]
