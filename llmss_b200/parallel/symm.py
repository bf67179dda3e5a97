"""Symmetric-memory workspace: one cudaMalloc'd region per rank, mapped into every peer with
CUDA IPC once at start-up (the handle exchange is the only thing that uses the c10d group).

This is the B200-native replacement for "NCCL communicator bring-up" (reference dist.py:44-73)
on the hot path: afterwards the kernels in csrc/comm.cu and the fused GEMM epilogue in
csrc/gemm_tcgen05.cu write straight into peer HBM over NVLink 5 / NVSwitch and synchronise
with system-scope flags, with no host or NCCL involvement (CUDA-graph capturable).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from ..ops import _lib

_DT = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2}


class SymmetricWorkspace:
    def __init__(self, group, device: torch.device, max_bytes: int = 32 << 20, lamport_bytes: int = 16 << 20):
        lib = _lib.load()
        self.group = group
        self.rank = group.rank()
        self.nranks = group.size()
        self.device = device
        self.buf_bytes = int(max_bytes)          # flag-synchronised bulk area (two-shot all-reduce, all-gather)
        self.lam_bytes = int(lamport_bytes)      # self-validating low-latency area (one-shot AR, fused GEMM+AR)
        with torch.cuda.device(device):
            base = C.c_void_p()
            _lib.check(lib.llmss_symm_alloc(self.buf_bytes, self.lam_bytes, C.byref(base)), "symm_alloc")
            self._local = base
            hbuf = C.create_string_buffer(64)
            _lib.check(lib.llmss_ipc_get_handle(base, hbuf), "ipc_get_handle")
            handles = [None] * self.nranks
            dist.all_gather_object(handles, hbuf.raw, group=group)
            bases = (C.c_void_p * 8)()
            self._opened = []
            for r in range(self.nranks):
                if r == self.rank:
                    bases[r] = base.value
                else:
                    p = C.c_void_p()
                    _lib.check(lib.llmss_ipc_open_handle(handles[r], C.byref(p)), f"ipc_open_handle(rank {r})")
                    bases[r] = p.value
                    self._opened.append(p)
            h = C.c_void_p()
            _lib.check(lib.llmss_symm_create(self.rank, self.nranks, bases, self.buf_bytes, self.lam_bytes, C.byref(h)),
                       "symm_create")
            self.handle = h
        # everyone has mapped everyone before the first kernel touches a peer
        dist.barrier(group=group, device_ids=[device.index])

    # ------------------------------------------------------------------ capability checks
    def can_all_reduce(self, x: torch.Tensor) -> bool:
        nbytes = x.numel() * x.element_size()
        if x.dtype not in _DT or not x.is_contiguous() or nbytes % 16 or x.data_ptr() % 16:
            return False
        if nbytes <= (1 << 20) and nbytes * self.nranks <= self.lam_bytes:
            return True  # one-shot (Lamport area)
        if (nbytes // 16) % self.nranks != 0:
            return nbytes * self.nranks <= self.lam_bytes
        return nbytes * 2 <= self.buf_bytes  # two-shot (bulk area)

    def can_all_gather(self, x: torch.Tensor) -> bool:
        nbytes = x.numel() * x.element_size()
        return x.is_contiguous() and nbytes % 16 == 0 and x.data_ptr() % 16 == 0 and nbytes * self.nranks <= self.buf_bytes

    # ------------------------------------------------------------------ collectives
    def all_reduce_(self, x: torch.Tensor, algo: int = 0) -> torch.Tensor:
        lib = _lib.load()
        s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        p = C.c_void_p(x.data_ptr())
        _lib.check(lib.llmss_allreduce(self.handle, p, p, x.numel(), _DT[x.dtype], algo, s), "llmss_allreduce")
        _lib.LAUNCHES["count"] += 1
        return x

    def all_gather(self, x: torch.Tensor) -> torch.Tensor:
        lib = _lib.load()
        out = torch.empty((self.nranks,) + tuple(x.shape), dtype=x.dtype, device=x.device)
        s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        _lib.check(lib.llmss_allgather(self.handle, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()),
                                       x.numel() * x.element_size(), s), "llmss_allgather")
        _lib.LAUNCHES["count"] += 1
        return out

    def all_gather_cols(self, x: torch.Tensor) -> torch.Tensor:
        """cat(all_gather(x), dim=-1) for a contiguous 2-D x, written in place by the gather kernel."""
        lib = _lib.load()
        rows, cols = x.shape
        out = torch.empty(rows, cols * self.nranks, dtype=x.dtype, device=x.device)
        s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        _lib.check(lib.llmss_allgather_cols(self.handle, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), rows,
                                            cols * x.element_size(), s), "llmss_allgather_cols")
        _lib.LAUNCHES["count"] += 1
        return out

    def epoch(self) -> int:
        v = C.c_uint32()
        _lib.check(_lib.load().llmss_symm_epoch(self.handle, C.byref(v)), "symm_epoch")
        return v.value
