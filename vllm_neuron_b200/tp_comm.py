"""Peer-memory plumbing for the tensor-parallel exchange step (b200_allreduce_rmsnorm).

torch.distributed's symmetric memory is used ONLY to obtain peer-mapped device memory (one region per rank with the same
layout, every region mapped into every process); the exchange itself is our kernel (csrc/tp_allreduce.cu), not NCCL.
Region layout (bytes):  [ partial A : rows*hidden bf16 | partial B : rows*hidden bf16 | flags : rows*8 uint32 ].
The row-parallel GEMMs (o_proj, down_proj) write their outputs straight into partial A / B of the local region.
Where the region also has a MULTICAST mapping (NVSwitch), the exchange kernel sums the ranks inside the switch
(multimem.ld_reduce on `multicast_ptr`, b200_allreduce_rmsnorm_nvls); otherwise it pulls every peer's rows.
"""
from __future__ import annotations

import ctypes as C

import torch


class TpComm:
    FLAG_SLOTS = 8  # = AR_MAX_TP in csrc/tp_allreduce.cu

    def __init__(self, group, rank: int, world: int, rows: int, hidden: int, device, use_nvls: bool | None = None):
        import os

        import torch.distributed._symmetric_memory as symm_mem

        self.rank, self.world, self.rows, self.hidden = rank, world, rows, hidden
        part_bytes = rows * hidden * 2
        part_bytes = (part_bytes + 255) // 256 * 256
        flag_bytes = rows * self.FLAG_SLOTS * 4
        # tail of the region: the greedy-merge exchange -- [rows] fp32 maxima | [rows] int64 indices | [8] uint32 flags
        flag_bytes = (flag_bytes + 255) // 256 * 256
        merge_off = 2 * part_bytes + flag_bytes
        merge_bytes = (rows * 4 + 255) // 256 * 256 + rows * 8 + 256
        self._region = symm_mem.empty((merge_off + merge_bytes,), dtype=torch.uint8, device=device)
        self._region.zero_()
        self._hdl = symm_mem.rendezvous(self._region, group)
        base = [int(p) for p in self._hdl.buffer_ptrs]
        assert len(base) == world and base[rank] == self._region.data_ptr()
        self.partials = [self._region[i * part_bytes: i * part_bytes + rows * hidden * 2].view(torch.bfloat16).view(rows, hidden)
                         for i in range(2)]
        arr = C.c_void_p * world
        self._peer_bufs = [arr(*[b + i * part_bytes for b in base]) for i in range(2)]
        self._peer_flags = arr(*[b + 2 * part_bytes for b in base])
        # NVLS: the region's multicast mapping (NVSwitch boxes; 0 where the fabric has none).  B200_AR_NVLS=0 keeps the pull.
        mc = int(getattr(self._hdl, "multicast_ptr", 0) or 0)
        if use_nvls is None:
            # measured (profiles/r02_tp_exchange_nvls_v1.md): no gain at TP=2 (the pull moves the same bytes and is the
            # bit-exact rank-order sum); +5..7 % whole-model tokens/s at TP=8, where the pull's 7x ingress contends
            use_nvls = os.environ.get("B200_AR_NVLS", "1" if world >= 4 else "0") != "0"
        self.nvls = bool(mc) and use_nvls and world > 1
        self._mc_bufs = [mc + i * part_bytes if self.nvls else 0 for i in range(2)]
        self.epoch = torch.zeros((rows,), dtype=torch.int32, device=device)
        v_off = merge_off
        i_off = v_off + (rows * 4 + 255) // 256 * 256
        f_off = i_off + rows * 8
        self.merge_vals = self._region[v_off: v_off + rows * 4].view(torch.float32)
        self.merge_idxs = self._region[i_off: i_off + rows * 8].view(torch.int64)
        self._peer_mvals = arr(*[b + v_off for b in base])
        self._peer_midxs = arr(*[b + i_off for b in base])
        self._peer_mflags = arr(*[b + f_off for b in base])
        self.merge_epoch = torch.zeros((1,), dtype=torch.int32, device=device)
        self._next = 0
        torch.cuda.synchronize(device)
        self._hdl.barrier()  # every rank has zeroed its flags before anyone signals

    def next_partial(self, t: int):
        """The buffer the next row-parallel GEMM must write ([t, hidden] view); alternates A / B."""
        i = self._next
        return i, self.partials[i][:t]

    def merge_argmax(self, b: int):
        """Global greedy tokens [b] from the (max, index) pairs every rank's argmax wrote into merge_vals / merge_idxs."""
        from . import ops

        return ops.tp_argmax_merge(self._peer_mvals, self._peer_midxs, self._peer_mflags, self.merge_epoch, self.rank,
                                   self.world, b)

    def exchange(self, i: int, t: int, norm_w, eps: float, quant):
        from . import ops

        assert i == self._next, "exchange sites must alternate between the two partial buffers"
        self._next ^= 1
        return ops.allreduce_rmsnorm(self._peer_bufs[i], self._peer_flags, self.epoch, self.rank, self.world, t, self.hidden,
                                     norm_w, eps, quant, mc_buf=self._mc_bufs[i])
