"""Symmetric-memory workspace: one cudaMalloc'd region per rank, mapped into every peer with
CUDA IPC once at start-up (the handle exchange is the only thing that uses the c10d group).

This is the B200-native replacement for "NCCL communicator bring-up" (reference dist.py:44-73)
on the hot path: afterwards the kernels in csrc/comm.cu and the fused GEMM epilogue in
csrc/gemm_tcgen05.cu write straight into peer HBM over NVLink 5 / NVSwitch and synchronise
with system-scope flags, with no host or NCCL involvement (CUDA-graph capturable).
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.distributed as dist

from ..ops import _lib

_DT = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2}


class SymmetricWorkspace:
    """One symmetric region per rank, mapped into every peer, optionally bound to an NVLS multicast object.

    Bring-up order (``LLMSS_SYMM=auto|vmm|ipc``, ``LLMSS_NVLS=0`` keeps the heap P2P-only):

    1. VMM heap (csrc/vmm.cu): ``cuMemCreate`` with a POSIX-fd handle, descriptors passed between the ranks
       over abstract Unix sockets (SCM_RIGHTS), every peer's allocation mapped locally;
    2. NVLS: rank 0 creates the multicast object, every rank adds its device and binds its allocation,
       and maps the object -- ``mc_base``: a store there lands on every GPU, ``multimem.ld_reduce`` sums
       all copies inside the switch;
    3. fall-back when (1) is refused on any rank: ``cudaMalloc`` + CUDA IPC (P2P only).

    Every decision is taken collectively (a step counts as failed if it failed on any rank), so all ranks
    always end up with the same kind of workspace.
    """

    def __init__(self, group, device: torch.device, max_bytes: int = 32 << 20, lamport_bytes: int = 16 << 20):
        lib = _lib.load()
        self.group = group
        self.rank = group.rank()
        self.nranks = group.size()
        self.device = device
        self.buf_bytes = int(max_bytes)          # flag-synchronised bulk area (two-shot all-reduce, all-gather, NVLS staging)
        self.lam_bytes = int(lamport_bytes)      # self-validating low-latency area (one-shot AR, fused GEMM+AR)
        self.mode = None                         # "vmm" | "ipc"
        self.has_multicast = False
        self.bringup_log = []
        self._region = None
        self._opened = []
        self._imported = []
        want = os.environ.get("LLMSS_SYMM", "auto")
        with torch.cuda.device(device):
            bases = mc_ptr = None
            if want in ("auto", "vmm"):
                try:
                    bases, mc_ptr = self._bringup_vmm(lib)
                except Exception as e:  # noqa: BLE001 -- collective failures are already agreed on inside
                    self.bringup_log.append(f"vmm heap unavailable: {e}")
                    bases = None
                if bases is None and want == "vmm":
                    raise RuntimeError("LLMSS_SYMM=vmm but the VMM heap could not be set up: " + "; ".join(self.bringup_log))
            if bases is None:
                bases = self._bringup_ipc(lib)
            h = C.c_void_p()
            _lib.check(lib.llmss_symm_create2(self.rank, self.nranks, bases, mc_ptr, self.buf_bytes, self.lam_bytes,
                                              C.byref(h)), "symm_create")
            self.handle = h
            self.has_multicast = bool(lib.llmss_symm_has_multicast(h))
        # everyone has mapped everyone before the first kernel touches a peer
        dist.barrier(group=group, device_ids=[device.index])

    # ------------------------------------------------------------------ bring-up
    def _all_ok(self, ok: bool, what: str) -> bool:
        flags = [None] * self.nranks
        dist.all_gather_object(flags, bool(ok), group=self.group)
        if not all(flags):
            self.bringup_log.append(f"{what}: failed on rank(s) {[i for i, f in enumerate(flags) if not f]}")
            return False
        return True

    def _bringup_ipc(self, lib):
        base = C.c_void_p()
        _lib.check(lib.llmss_symm_alloc(self.buf_bytes, self.lam_bytes, C.byref(base)), "symm_alloc")
        self._local = base
        hbuf = C.create_string_buffer(64)
        _lib.check(lib.llmss_ipc_get_handle(base, hbuf), "ipc_get_handle")
        handles = [None] * self.nranks
        dist.all_gather_object(handles, hbuf.raw, group=self.group)
        bases = (C.c_void_p * 8)()
        for r in range(self.nranks):
            if r == self.rank:
                bases[r] = base.value
            else:
                p = C.c_void_p()
                _lib.check(lib.llmss_ipc_open_handle(handles[r], C.byref(p)), f"ipc_open_handle(rank {r})")
                bases[r] = p.value
                self._opened.append(p)
        self.mode = "ipc"
        return bases

    def _bringup_vmm(self, lib):
        caps = C.c_int(0)
        rc = lib.llmss_vmm_caps(C.byref(caps))
        if not self._all_ok(rc == 0 and (caps.value & 1), "vmm capability (posix-fd handles)"):
            return None, None
        want_mc = os.environ.get("LLMSS_NVLS", "1") != "0" and self.nranks > 1
        mc_caps = self._all_ok(bool(caps.value & 2), "multicast capability") if want_mc else False
        region_bytes = lib.llmss_symm_region_bytes(self.buf_bytes, self.lam_bytes)
        padded = C.c_size_t(0)
        rc = lib.llmss_vmm_padded_size(region_bytes, self.nranks, 1 if mc_caps else 0, C.byref(padded))
        sizes = [None] * self.nranks
        dist.all_gather_object(sizes, (rc, padded.value), group=self.group)
        if any(s[0] != 0 for s in sizes) or len({s[1] for s in sizes}) != 1:
            self.bringup_log.append(f"vmm padded size disagrees / failed: {sizes}")
            return None, None
        padded = padded.value
        region = C.c_void_p()
        rc = lib.llmss_vmm_alloc(padded, C.byref(region))
        err = _lib.last_error() if rc != 0 else ""
        if not self._all_ok(rc == 0, f"cuMemCreate ({err})"):
            if rc == 0:
                lib.llmss_vmm_free(region)
            return None, None
        self._region = region
        base = lib.llmss_vmm_ptr(region)
        _lib.check(lib.llmss_symm_init_region(C.c_void_p(base), self.buf_bytes, self.lam_bytes), "symm_init_region")

        # ---- descriptor exchange over abstract unix datagram sockets
        import uuid

        tok = [uuid.uuid4().hex[:12] if self.rank == 0 else None]
        dist.broadcast_object_list(tok, src=0, group=self.group)
        name = lambda r: f"llmss-{tok[0]}-{r}".encode()  # noqa: E731
        sock = C.c_int(-1)
        rc = lib.llmss_fdx_open(name(self.rank), C.byref(sock))
        if not self._all_ok(rc == 0, "fd-exchange socket"):
            lib.llmss_fdx_close(sock)
            self._free_region(lib)
            return None, None
        ok = True
        bases = (C.c_void_p * 8)()
        bases[self.rank] = base
        try:
            for p in range(1, self.nranks):
                peer = (self.rank + p) % self.nranks
                fd = C.c_int(-1)
                _lib.check(lib.llmss_vmm_export_fd(region, C.byref(fd)), "vmm_export_fd")
                try:
                    _lib.check(lib.llmss_fdx_send(sock, name(peer), self.rank, fd), f"fdx_send(rank {peer})")
                finally:
                    lib.llmss_fd_close(fd)
            for _ in range(self.nranks - 1):
                tag, fd = C.c_int(-1), C.c_int(-1)
                _lib.check(lib.llmss_fdx_recv(sock, 30000, C.byref(tag), C.byref(fd)), "fdx_recv")
                ptr = C.c_void_p()
                _lib.check(lib.llmss_vmm_import(fd, padded, C.byref(ptr)), f"vmm_import(rank {tag.value})")
                bases[tag.value] = ptr.value
                self._imported.append((ptr, padded))
        except Exception as e:  # noqa: BLE001
            self.bringup_log.append(f"rank {self.rank}: {e}")
            ok = False
        if not self._all_ok(ok, "peer mapping over VMM handles"):
            lib.llmss_fdx_close(sock)
            self._free_region(lib)
            return None, None
        self.mode = "vmm"

        # ---- NVLS multicast (optional)
        mc_ptr = None
        if mc_caps:
            mc = C.c_ulonglong(0)
            ok = True
            try:
                if self.rank == 0:
                    fd = C.c_int(-1)
                    _lib.check(lib.llmss_mc_create(self.nranks, padded, C.byref(mc), C.byref(fd)), "mc_create")
                    try:
                        for peer in range(1, self.nranks):
                            _lib.check(lib.llmss_fdx_send(sock, name(peer), 1000, fd), f"fdx_send(mc, rank {peer})")
                    finally:
                        lib.llmss_fd_close(fd)
            except Exception as e:  # noqa: BLE001
                self.bringup_log.append(f"rank 0: {e}")
                ok = False
            # rank 0's outcome decides whether the peers wait for a descriptor at all
            st = [ok if self.rank == 0 else None]
            dist.broadcast_object_list(st, src=0, group=self.group)
            if st[0]:
                try:
                    if self.rank != 0:
                        tag, fd = C.c_int(-1), C.c_int(-1)
                        _lib.check(lib.llmss_fdx_recv(sock, 30000, C.byref(tag), C.byref(fd)), "fdx_recv(mc)")
                        _lib.check(lib.llmss_mc_import(fd, C.byref(mc)), "mc_import")
                    _lib.check(lib.llmss_mc_add_device(mc), "mc_add_device")
                except Exception as e:  # noqa: BLE001
                    self.bringup_log.append(f"rank {self.rank}: {e}")
                    ok = False
                if self._all_ok(ok, "multicast add-device"):  # doubles as the barrier before any bind
                    rc = lib.llmss_mc_bind_and_map(mc, region)
                    err = _lib.last_error() if rc != 0 else ""
                    if self._all_ok(rc == 0, f"multicast bind ({err})"):
                        mc_ptr = C.c_void_p(lib.llmss_vmm_mc_ptr(region))
                    else:
                        self.bringup_log.append("NVLS disabled: bind failed")
                elif mc.value:
                    lib.llmss_mc_release(mc)
            else:
                self.bringup_log.append("NVLS disabled: cuMulticastCreate refused")
        lib.llmss_fdx_close(sock)
        return bases, mc_ptr

    def _free_region(self, lib):
        for ptr, padded in self._imported:
            lib.llmss_vmm_unmap(ptr, padded)
        self._imported = []
        if self._region is not None:
            lib.llmss_vmm_free(self._region)
            self._region = None

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
