"""ctypes binding of libllmss_b200.so (the C ABI in include/llmss_b200.h).

The library is loaded lazily.  On a machine with a GPU a missing / stale library is a hard
error (no silent PyTorch fallback on the CUDA path); ``LLMSS_B200_BUILD=1`` (default) lets
the loader run the in-tree nvcc build first.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_LIB = None
LOADED_FROM = None

c_void_p, c_int, c_int64, c_float, c_size_t = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t


class GemmPlan(C.Structure):
    _fields_ = [("bn", c_int), ("tiles", c_int), ("kb_total", c_int), ("grid", c_int), ("split", c_int),
                ("sk_ws_floats", c_int64), ("sk_counters", c_int64), ("ar_bytes", c_int64)]


_SIGS = {
    "llmss_last_error": (C.c_char_p, []),
    "llmss_version": (c_int, []),
    "llmss_symm_alloc": (c_int, [c_size_t, c_size_t, C.POINTER(c_void_p)]),
    "llmss_symm_free": (c_int, [c_void_p]),
    "llmss_ipc_get_handle": (c_int, [c_void_p, C.c_char_p]),
    "llmss_ipc_open_handle": (c_int, [C.c_char_p, C.POINTER(c_void_p)]),
    "llmss_ipc_close": (c_int, [c_void_p]),
    "llmss_symm_region_bytes": (c_size_t, [c_size_t, c_size_t]),
    "llmss_symm_create": (c_int, [c_int, c_int, C.POINTER(c_void_p), c_size_t, c_size_t, C.POINTER(c_void_p)]),
    "llmss_symm_destroy": (c_int, [c_void_p]),
    "llmss_symm_init_region": (c_int, [c_void_p, c_size_t, c_size_t]),
    "llmss_symm_create2": (c_int, [c_int, c_int, C.POINTER(c_void_p), c_void_p, c_size_t, c_size_t, C.POINTER(c_void_p)]),
    "llmss_symm_has_multicast": (c_int, [c_void_p]),
    "llmss_debug_wait_flag": (c_int, [c_void_p, C.c_uint32, c_void_p]),
    "llmss_vmm_caps": (c_int, [C.POINTER(c_int)]),
    "llmss_vmm_padded_size": (c_int, [c_size_t, c_int, c_int, C.POINTER(c_size_t)]),
    "llmss_vmm_alloc": (c_int, [c_size_t, C.POINTER(c_void_p)]),
    "llmss_vmm_ptr": (c_void_p, [c_void_p]),
    "llmss_vmm_mc_ptr": (c_void_p, [c_void_p]),
    "llmss_vmm_bytes": (c_size_t, [c_void_p]),
    "llmss_vmm_export_fd": (c_int, [c_void_p, C.POINTER(c_int)]),
    "llmss_vmm_import": (c_int, [c_int, c_size_t, C.POINTER(c_void_p)]),
    "llmss_vmm_unmap": (c_int, [c_void_p, c_size_t]),
    "llmss_vmm_free": (c_int, [c_void_p]),
    "llmss_mc_create": (c_int, [c_int, c_size_t, C.POINTER(C.c_ulonglong), C.POINTER(c_int)]),
    "llmss_mc_import": (c_int, [c_int, C.POINTER(C.c_ulonglong)]),
    "llmss_mc_add_device": (c_int, [C.c_ulonglong]),
    "llmss_mc_bind_and_map": (c_int, [C.c_ulonglong, c_void_p]),
    "llmss_mc_release": (c_int, [C.c_ulonglong]),
    "llmss_fdx_open": (c_int, [C.c_char_p, C.POINTER(c_int)]),
    "llmss_fdx_send": (c_int, [c_int, C.c_char_p, c_int, c_int]),
    "llmss_fdx_recv": (c_int, [c_int, c_int, C.POINTER(c_int), C.POINTER(c_int)]),
    "llmss_fdx_close": (c_int, [c_int]),
    "llmss_fd_close": (c_int, [c_int]),
    "llmss_symm_epoch": (c_int, [c_void_p, C.POINTER(C.c_uint32)]),
    "llmss_allreduce": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    "llmss_allgather": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "llmss_argmax_local": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_void_p]),
    "llmss_argmax_merge": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "llmss_vp_record_floats": (c_int, [c_int]),
    "llmss_vp_summary": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                 c_void_p]),
    "llmss_vp_multinomial_pick": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_int, c_int, c_int, c_void_p,
                                          c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "llmss_vp_multinomial_merge": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "llmss_vp_merge_sample": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_void_p, c_void_p]),
    "llmss_allgather_cols": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_void_p]),
    "llmss_gemm_set_debug": (c_int, [c_void_p]),
    "llmss_gemm_set_pair": (c_int, [c_int]),
    "llmss_gemm_plan_query": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int, C.POINTER(GemmPlan)]),
    "llmss_gemm": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_void_p, c_int64,
                           c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                           c_int, c_void_p]),
    "llmss_gemm_ex": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_void_p, c_int64,
                              c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                              c_int, c_int, c_void_p, c_void_p]),
    "llmss_norm_residual": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                    c_int, c_int64, c_int64, c_int64, c_int64, c_float, c_int, c_int, c_void_p]),
    "llmss_rope_kv_write": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int, c_int,
                                    c_int, c_int, c_int, c_float, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "llmss_attn_decode_scratch_floats": (c_int64, [c_int, c_int, c_int, c_int]),
    "llmss_attn_decode_pick_splits": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "llmss_attn_decode_tc_pick_splits": (c_int, [c_int, c_int, c_int, c_int]),
    "llmss_attn_decode_tc_balanced_scratch_floats": (c_int64, [c_int, c_int, c_int, c_int, c_int]),
    "llmss_attn_decode_tc": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                     c_int64, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_float, c_int,
                                     c_int, c_void_p]),
    "llmss_attn_decode": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                  c_int64, c_void_p, c_int, c_int, c_int, c_int, c_int, c_float, c_int, c_int,
                                  c_void_p]),
    "llmss_embed": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                            c_int64, c_int, c_void_p]),
    "llmss_sample": (c_int, [c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                             c_void_p, c_void_p, c_void_p]),
}
_OPTIONAL = {
    "llmss_attn_prefill_paged": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int64,
                                         c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_float, c_int,
                                         c_void_p]),
    "llmss_attn_prefill": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64,
                                   c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_float, c_int, c_void_p]),
}


def lib_path() -> Path:
    return Path(__file__).resolve().parents[1] / "_C" / "libllmss_b200.so"


def load(build_if_missing: bool = True):
    global _LIB, LOADED_FROM
    if _LIB is not None:
        return _LIB
    path = lib_path()
    override = os.environ.get("LLMSS_B200_LIB")  # A/B a previously built library (profiling aid)
    if override:
        from pathlib import Path

        path, build_if_missing = Path(override), False
    if build_if_missing and os.environ.get("LLMSS_B200_BUILD", "1") == "1":
        try:
            from .build import build_library

            build_library()
        except Exception as e:  # nvcc may be absent on a deployment box: fall through to dlopen
            if not path.exists():
                raise RuntimeError(f"libllmss_b200.so is missing and could not be built: {e}") from e
    if not path.exists():
        raise RuntimeError(f"{path} not found; run `python -m llmss_b200.ops.build`")
    lib = C.CDLL(str(path))
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    for name, (res, args) in _OPTIONAL.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
    _LIB, LOADED_FROM = lib, str(path)
    return lib


def has(name: str) -> bool:
    return hasattr(load(), name)


class KernelError(RuntimeError):
    pass


def last_error() -> str:
    msg = load().llmss_last_error()
    return msg.decode() if msg else ""


def check(rc: int, what: str):
    if rc != 0:
        msg = load().llmss_last_error()
        raise KernelError(f"{what} failed (rc={rc}): {msg.decode() if msg else '?'}")


# number of kernel launches issued through this binding (bench.py reports it)
LAUNCHES = {"count": 0}
