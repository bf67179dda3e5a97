"""Collective front door used by every layer of the framework.

The reference calls `torch.distributed.all_reduce / all_gather / broadcast` inline at each
call site (`layers.py:125,133,178,213`, `generate.py:85,144`).
Here every call site goes through a :class:`Communicator`, which picks per message:

  * world_size == 1 / `FakeGroup`         -> no-op (fixes SURVEY Q2)
  * CUDA + symmetric-memory workspace       -> hand-written one-shot / two-shot P2P kernels
                                              over NVLink / NVSwitch multicast (csrc/comm.cu), CUDA-graph safe
  * otherwise (CPU gloo, or message too big for the workspace) -> torch.distributed

The fused GEMM->all-reduce kernel does not go through here at all: it owns the same
workspace and performs the exchange from its epilogue (csrc/gemm_tcgen05.cu).
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

from ..utils.dist import FakeGroup, is_fake


class Communicator:
    def __init__(self, process_group=None, device: Optional[torch.device] = None):
        self.group = process_group
        self.fake = is_fake(process_group)
        if process_group is None:
            self.rank, self.size = 0, 1
        else:
            self.rank, self.size = process_group.rank(), process_group.size()
        self.device = device
        self.symm = None  # SymmetricWorkspace, attached by attach_symmetric()
        self.stats = {"all_reduce": 0, "all_gather": 0, "broadcast": 0, "symm_all_reduce": 0}

    # ------------------------------------------------------------------ setup
    def attach_symmetric(self, max_bytes: int = 32 << 20, lamport_bytes: int = 16 << 20):
        """Allocate + exchange the NVLink peer-mapped workspace (CUDA, size > 1 only)."""
        if self.size == 1 or self.fake or self.device is None or self.device.type != "cuda":
            return None
        from .symm import SymmetricWorkspace

        self.symm = SymmetricWorkspace(self.group, self.device, max_bytes=max_bytes, lamport_bytes=lamport_bytes)
        return self.symm

    @property
    def single(self) -> bool:
        return self.size == 1 or self.fake

    # ------------------------------------------------------------- collectives
    def all_reduce(self, x: torch.Tensor) -> torch.Tensor:
        """In-place SUM over the TP group (reference N1/N2)."""
        self.stats["all_reduce"] += 1
        if self.single:
            return x
        if self.symm is not None and x.is_cuda and self.symm.can_all_reduce(x):
            self.stats["symm_all_reduce"] += 1
            return self.symm.all_reduce_(x)
        dist.all_reduce(x, group=self.group)
        return x

    def all_gather_cat(self, x: torch.Tensor, dim: int = -1) -> torch.Tensor:
        """Gather equally-shaped shards and concatenate along ``dim`` (reference N3)."""
        self.stats["all_gather"] += 1
        if self.size == 1:
            return x
        if self.fake:  # DEBUG=1 shard inspection: every slot is the local shard
            return torch.cat([x] * self.size, dim=dim)
        if self.symm is not None and x.is_cuda and self.symm.can_all_gather(x):
            if x.dim() == 2 and dim in (-1, 1) and (x.shape[1] * x.element_size()) % 16 == 0:
                return self.symm.all_gather_cols(x)  # concatenated in place, no torch.cat pass
            parts = self.symm.all_gather(x)  # [t, *x.shape]
            parts = list(parts.unbind(0))
        else:
            x = x.contiguous()
            parts = [torch.empty_like(x) for _ in range(self.size)]
            dist.all_gather(parts, x, group=self.group)
        return torch.cat(parts, dim=dim)

    def broadcast(self, x: torch.Tensor, src: int = 0) -> torch.Tensor:
        self.stats["broadcast"] += 1
        if self.single:
            return x
        dist.broadcast(x, src=src, group=self.group)
        return x

    def broadcast_object(self, obj, src: int = 0):
        if self.single:
            return obj
        box = [obj if self.rank == src else None]
        dist.broadcast_object_list(box, src=src, group=self.group)
        return box[0]

    def barrier(self):
        if self.single:
            return
        if self.device is not None and self.device.type == "cuda":
            dist.barrier(group=self.group, device_ids=[self.device.index])
        else:
            dist.barrier(group=self.group)

    def max_over_ranks(self, value: float) -> float:
        if self.single:
            return value
        dev = self.device if self.device is not None else torch.device("cpu")
        t = torch.tensor([value], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return float(t.item())


_COMMS: dict = {}


def comm_for(process_group, device=None) -> Communicator:
    """One Communicator per (group, device); created on first use."""
    dev = torch.device(device) if device is not None else None
    key = (id(process_group), str(dev))
    c = _COMMS.get(key)
    if c is None:
        # re-use a communicator created for the same group on any device (keeps the workspace)
        for (gid, _), cc in _COMMS.items():
            if gid == id(process_group) and (dev is None or cc.device == dev or cc.device is None):
                if cc.device is None and dev is not None:
                    cc.device = dev
                return cc
        c = Communicator(process_group, dev)
        _COMMS[key] = c
    return c


def reset_comms():
    _COMMS.clear()
