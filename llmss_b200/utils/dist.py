"""Process-group bring-up.

Parity: reference `src/llmss/server/models/utils/dist.py:9-77`
(`initialize_torch_distributed`, `FakeGroup`, `FakeBarrier`).

Differences (deliberate):
  * `FakeGroup` here is a *complete* in-process group: besides the c10d duck-typed
    methods of the reference (`allreduce`/`allgather`/`barrier`/`size`/`rank`) it is
    understood by every collective helper in :mod:`llmss_b200.parallel.comm`, so
    world_size == 1 works end-to-end (the reference's `generate.py` cannot run at
    WORLD_SIZE=1 because it calls `dist.broadcast` without a default group).
  * One process per GPU, single node (same as the reference), device chosen from
    LOCAL_RANK when present.
"""
from __future__ import annotations

import os
from datetime import timedelta

import torch

from .logging import logger


class FakeBarrier:
    """Completed-work handle returned by :class:`FakeGroup` (ref dist.py:9-11)."""

    def wait(self, *args, **kwargs):
        return True

    def is_completed(self):
        return True


class FakeGroup:
    """Communication-free stand-in for a process group (ref dist.py:14-37).

    Used when world_size == 1, or when ``DEBUG=1`` so that one process can build
    and inspect "rank r of N" shards without peers (results are then numerically
    partial, exactly like the reference).
    """

    def __init__(self, rank: int, size: int):
        self._rank = rank
        self._size = size

    # c10d duck-typing -----------------------------------------------------
    def allreduce(self, *args, **kwargs):
        return FakeBarrier()

    def allgather(self, outputs, inputs, *args, **kwargs):
        # torch >= 2.1 passes an extra options argument; accept and ignore it
        # (the reference's signature breaks there).
        for out_list, inp in zip(outputs, inputs):
            for o in out_list:
                o.data = inp.data
        return FakeBarrier()

    def broadcast(self, *args, **kwargs):
        return FakeBarrier()

    def barrier(self, *args, **kwargs):
        return FakeBarrier()

    def size(self) -> int:
        return self._size

    def rank(self) -> int:
        return self._rank

    def __repr__(self):
        return f"FakeGroup(rank={self._rank}, size={self._size})"


def is_fake(group) -> bool:
    return group is None or isinstance(group, FakeGroup)


def initialize_torch_distributed(timeout_s: int = 60):
    """Return ``(process_group, rank, world_size)`` (ref dist.py:40-77).

    NCCL when CUDA is available (high-priority stream, ``timeout_s`` timeout),
    gloo otherwise; :class:`FakeGroup` for world_size == 1 or ``DEBUG=1``.
    """
    rank = int(os.getenv("RANK", "0"))
    world_size = int(os.getenv("WORLD_SIZE", "1"))

    options = None
    if torch.cuda.is_available():
        ndev = torch.cuda.device_count()
        assert world_size <= ndev, "Each process is one gpu"
        device = int(os.getenv("LOCAL_RANK", str(rank % ndev))) % ndev
        torch.cuda.set_device(device)
        backend = "nccl"
        try:
            from torch.distributed import ProcessGroupNCCL

            options = ProcessGroupNCCL.Options()
            options.is_high_priority_stream = True
            options._timeout = timedelta(seconds=timeout_s)
        except Exception:  # pragma: no cover - very old/new torch
            options = None
    else:
        backend = "gloo"

    if world_size == 1 or os.getenv("DEBUG", None) == "1":
        return FakeGroup(rank, world_size), rank, world_size

    if not torch.distributed.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        kwargs = dict(backend=backend, world_size=world_size, rank=rank,
                      timeout=timedelta(seconds=timeout_s))
        if options is not None:
            kwargs["pg_options"] = options
        if backend == "nccl":
            kwargs["device_id"] = torch.device("cuda", torch.cuda.current_device())
        torch.distributed.init_process_group(**kwargs)
    else:
        logger.warning("torch.distributed is already initialized.")

    return torch.distributed.group.WORLD, rank, world_size
