"""llmss_attn_decode_static: the split-KV decode attention entry point that allocates nothing.

Some hosts of this library cannot hand it scratch memory or let it allocate: a server that has
already given every byte of HBM to weights and KV pages, a decode loop that cannot take the
device-wide stall of cudaMalloc/cudaFree, or a C program with no allocator behind it.  For
them the split-KV partials live in an 8 MiB array inside the library.  This test drives that
entry point straight through the C ABI (ctypes, no Python wrapper) and checks that it returns
what the ordinary entry point (caller-provided scratch) returns -- twice in a row with different
shapes, because the second call finds the first call's leftovers in the shared array -- and
that a request beyond the built-in limit is refused with a readable error.

The file name sorts last on purpose: a run that stops at the first failure has then already
run every other test.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

DT = {torch.float16: 0, torch.bfloat16: 1}


def _lib():
    from llmss_b200.ops import _lib

    _lib.load()  # builds if needed; we then bind the raw symbols ourselves
    lib = C.CDLL(str(_lib.lib_path()))
    vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_float
    lib.llmss_last_error.restype = C.c_char_p
    lib.llmss_attn_decode_scratch_floats.restype = i64
    lib.llmss_attn_decode_scratch_floats.argtypes = [i32, i32, i32, i32]
    lib.llmss_attn_decode_static_max_floats.restype = i64
    lib.llmss_attn_decode.restype = i32
    lib.llmss_attn_decode.argtypes = [vp, i64, vp, vp, vp, i32, vp, vp, i64, vp, i32, i32, i32, i32, i32, f32, i32, i32, vp]
    lib.llmss_attn_decode_static.restype = i32
    lib.llmss_attn_decode_static.argtypes = [vp, i64, vp, vp, vp, i32, vp, vp, i64, i32, i32, i32, i32, i32, f32, i32, i32, vp]
    return lib


def _case(B, hq, hkv, d, lens, dtype, seed):
    torch.manual_seed(seed)
    bs = 16
    maxb = (max(lens) + bs - 1) // bs
    nb = 1 + B * maxb
    kc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=dtype)
    vc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=dtype)
    perm = (torch.randperm(nb - 1) + 1).tolist()
    bt = torch.tensor([[perm[i * maxb + j] for j in range(maxb)] for i in range(B)], dtype=torch.int32, device="cuda")
    sl = torch.tensor(lens, dtype=torch.int32, device="cuda")
    q = torch.randn(B, hq * d, device="cuda", dtype=dtype)
    return q, kc, vc, bt, sl, bs


def _run_both(lib, B, hq, hkv, d, lens, splits, dtype, seed):
    q, kc, vc, bt, sl, bs = _case(B, hq, hkv, d, lens, dtype, seed)
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    scale = d ** -0.5
    n = lib.llmss_attn_decode_scratch_floats(B, hq, d, splits)
    scratch = torch.empty(max(int(n), 1), dtype=torch.float32, device="cuda")
    want = torch.zeros(B, hq * d, device="cuda", dtype=dtype)
    got = torch.zeros(B, hq * d, device="cuda", dtype=dtype)
    rc = lib.llmss_attn_decode(p(q), q.stride(0), p(kc), p(vc), p(bt), bt.shape[1], p(sl), p(want), want.stride(0),
                               p(scratch), B, hq, hkv, d, bs, scale, splits, DT[dtype], stream)
    assert rc == 0, lib.llmss_last_error()
    rc = lib.llmss_attn_decode_static(p(q), q.stride(0), p(kc), p(vc), p(bt), bt.shape[1], p(sl), p(got), got.stride(0),
                                      B, hq, hkv, d, bs, scale, splits, DT[dtype], stream)
    assert rc == 0, lib.llmss_last_error()
    torch.cuda.synchronize()
    err = (got.float() - want.float()).abs()
    tol = 2e-2 + 2e-2 * want.float().abs()  # the tolerance test_attn_decode uses against the oracle
    assert int((err > tol).sum()) == 0, f"static vs. regular entry point: max err {err.max().item():.4g}"
    # both entry points launch the same deterministic kernels on the same inputs (fixed split partition, ordered
    # combine, no atomics); only the home of the partials differs, so the outputs must agree bit for bit
    assert torch.equal(got, want), f"static vs. regular entry point differ in bits: max err {err.max().item():.4g}"
    assert want.float().abs().sum().item() > 0


def test_static_entry_point_matches_regular_twice_in_a_row():
    lib = _lib()
    assert lib.llmss_attn_decode_static_max_floats() == 2 * 1024 * 1024
    # first call: large request that fills a good part of the built-in array
    _run_both(lib, B=24, hq=16, hkv=16, d=256, lens=[1000 + 7 * i for i in range(24)], splits=4,
              dtype=torch.bfloat16, seed=1)
    # second call straight after: different batch, heads, head_dim and split count, so its partial
    # slots overlap the first call's stale data at different offsets
    _run_both(lib, B=5, hq=8, hkv=2, d=128, lens=[1, 17, 300, 64, 1025], splits=8, dtype=torch.float16, seed=2)
    # no-split requests need no scratch and are always accepted
    _run_both(lib, B=3, hq=4, hkv=4, d=64, lens=[5, 90, 33], splits=1, dtype=torch.bfloat16, seed=3)


def test_static_entry_point_refuses_oversize_request():
    lib = _lib()
    B, hq, d, splits = 256, 64, 256, 16  # 256*64*16*258 floats = 270 MiB >> 8 MiB
    assert lib.llmss_attn_decode_scratch_floats(B, hq, d, splits) > lib.llmss_attn_decode_static_max_floats()
    x = torch.zeros(16, device="cuda")
    p = C.c_void_p(x.data_ptr())
    rc = lib.llmss_attn_decode_static(p, 8, p, p, p, 1, p, p, 8, B, hq, hq, d, 16, 1.0, splits, 1, None)
    assert rc != 0
    msg = lib.llmss_last_error().decode()
    assert "exceed" in msg and "8 MiB" in msg, msg


def test_static_entry_point_rejects_zero_kv_heads_without_crashing():
    """hkv = 0 used to reach `hq % hkv` on the host (integer division by zero kills the process); both entry
    points now answer with an error code and a message."""
    lib = _lib()
    q, kc, vc, bt, sl, bs = _case(2, 4, 4, 64, [5, 9], torch.bfloat16, 4)
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    out = torch.zeros(2, 4 * 64, device="cuda", dtype=torch.bfloat16)
    rc = lib.llmss_attn_decode_static(p(q), q.stride(0), p(kc), p(vc), p(bt), bt.shape[1], p(sl), p(out), out.stride(0),
                                      2, 4, 0, 64, bs, 0.125, 1, DT[torch.bfloat16], stream)
    assert rc != 0 and b"hkv" in lib.llmss_last_error()
    rc = lib.llmss_attn_decode(p(q), q.stride(0), p(kc), p(vc), p(bt), bt.shape[1], p(sl), p(out), out.stride(0), None,
                               2, 4, 0, 64, bs, 0.125, 1, DT[torch.bfloat16], stream)
    assert rc != 0 and b"hkv" in lib.llmss_last_error()
