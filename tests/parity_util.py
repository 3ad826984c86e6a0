"""Shared helpers for the parity tests (test infrastructure)."""
import torch


def logical_to_device_pool(pool_logical):
    """[nb, bs, Hkv, D] (oracle, token-major) -> [nb, Hkv, bs, D] (device layout)."""
    return pool_logical.permute(0, 2, 1, 3).contiguous()


def device_to_logical_pool(pool_dev):
    return pool_dev.permute(0, 2, 1, 3).contiguous()


def make_block_tables(b, ctx_lens, block_size, num_blocks, max_blocks, gen):
    """Random non-contiguous block ids from [1, num_blocks) (block 0 = null block), 0-padded rows."""
    need = [-(-int(c) // block_size) for c in ctx_lens]
    perm = (torch.randperm(num_blocks - 1, generator=gen) + 1).tolist()
    assert sum(need) <= len(perm)
    bt = torch.zeros((b, max_blocks), dtype=torch.int32)
    off = 0
    for i, n in enumerate(need):
        bt[i, :n] = torch.tensor(perm[off : off + n], dtype=torch.int32)
        off += n
    return bt


def rel_err(got, want):
    got, want = got.float(), want.float()
    return ((got - want).norm() / want.norm().clamp_min(1e-12)).item()


def max_rel_err(got, want, floor):
    got, want = got.float(), want.float()
    return ((got - want).abs() / want.abs().clamp_min(floor)).max().item()
