"""GPU numerics: every sm_100a kernel against the plain-PyTorch fp32 oracle of the same op."""
import math

import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from llmss_b200.ops import reference as R  # noqa: E402


def _cu():
    from llmss_b200.ops import cuda_impl

    return cuda_impl


def _close(a, b, atol, rtol, what=""):
    a, b = a.float(), b.float()
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    bad = (err > tol).sum().item()
    assert bad == 0, f"{what}: {bad}/{a.numel()} mismatches, max err {err.max().item():.4g}"


DT = [torch.bfloat16, torch.float16]


def test_library_is_loaded():
    from llmss_b200.ops import _lib

    lib = _lib.load()
    assert lib.llmss_version() >= 100
    assert "libllmss_b200.so" in _lib.LOADED_FROM


# ------------------------------------------------------------------ GEMM
GEMM_SHAPES = [
    # (M, N, K)   decode-like, stream-K splits, prefill-like, ragged edges
    (1, 128, 64), (1, 4096, 4096), (8, 512, 1024), (16, 384, 256), (17, 256, 512), (32, 4096, 1024),
    (64, 12288, 4096), (64, 4096, 2560), (100, 1000, 328), (128, 2048, 2048), (200, 768, 3072),
    (256, 1024, 1024), (300, 640, 192), (1024, 4096, 4096), (4096, 512, 256),
    (64, 28672, 4096), (64, 4096, 20480),   # the two GEMMs of the flagship decode step (GPT-J-6B, batch 64, TP=1)
]


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_plain(M, N, K, dtype):
    cu = _cu()
    torch.manual_seed(M * 7 + N * 3 + K)
    x = torch.randn(M, K, device="cuda", dtype=dtype)
    w = torch.randn(N, K, device="cuda", dtype=dtype) / math.sqrt(K)
    y = cu.linear(x, w)
    torch.cuda.synchronize()
    ref = R.linear(x, w, out_dtype=torch.float32)
    _close(y, ref, 3e-2 if dtype == torch.bfloat16 else 6e-3, 2e-2, f"gemm {M}x{N}x{K}")


def test_gemm_repeatable_and_counter_reset():
    """stream-K counters must return to zero: the same launch twice gives identical bits."""
    cu = _cu()
    x = torch.randn(48, 4096, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(4096, 4096, device="cuda", dtype=torch.bfloat16) / 64
    a = cu.linear(x, w).clone()
    b = cu.linear(x, w).clone()
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    assert int(cu.arena(x.device).sk_counters.abs().sum().item()) == 0


@pytest.mark.parametrize("act", ["gelu_new", "gelu", "relu", "silu"])
def test_gemm_bias_act_residual(act):
    cu = _cu()
    M, N, K = 40, 1536, 512
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / math.sqrt(K)
    b = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    res = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
    a = R.ACT_IDS[act]
    y = cu.linear(x, w, b, act=a, residual=res)
    _close(y, R.linear(x, w, b, act=a, residual=res, out_dtype=torch.float32), 4e-2, 2e-2, act)
    y2 = cu.linear(x, w, b, act=a, act_start=512)
    _close(y2, R.linear(x, w, b, act=a, act_start=512, out_dtype=torch.float32), 4e-2, 2e-2, act + " act_start")


def test_gemm_gated_and_f32_out_and_strided_input():
    cu = _cu()
    M, N, K = 24, 1024, 768
    big = torch.randn(M, K + 256, device="cuda", dtype=torch.float16)
    x = big[:, 256:]  # row stride != K
    w = torch.randn(N, K, device="cuda", dtype=torch.float16) / math.sqrt(K)
    y = cu.linear(x, w, act=R.ACT_SILU, gated=True)
    assert y.shape == (M, N // 2)
    _close(y, R.linear(x, w, act=R.ACT_SILU, gated=True, out_dtype=torch.float32), 1e-2, 2e-2, "gated")
    z = cu.linear(x, w, out_dtype=torch.float32)
    assert z.dtype == torch.float32
    _close(z, R.linear(x, w, out_dtype=torch.float32), 5e-3, 1e-2, "f32 out")


@pytest.mark.parametrize("mode", [8, 1, 2, 4])
def test_gemm_cta_pair_kernel(mode):
    """tcgen05 cta_group::2 kernel (two CTAs share one 256 x 256 MMA): the automatic choice (persistent pairs from 40
    pair tiles upwards) and the forced schedules (1 = persistent, 2 / 4 = split-K over 2 / 4 pairs per cluster), every
    epilogue flavour, ragged rows / odd feature-tile counts / K tails, against the fp32 oracle and the single-CTA kernels."""
    from llmss_b200.ops import _lib

    cu = _cu()
    lib = _lib.load()
    cases = [(256, 12288, 512, torch.bfloat16), (700, 10240 + 128, 320, torch.bfloat16), (129, 10496, 1088, torch.float16)]
    if mode != 8:
        cases += [(256, 512, 512, torch.bfloat16), (200, 4224, 4160, torch.bfloat16)]
    else:
        cases += [(256, 6144, 512, torch.bfloat16)]     # 24 pair tiles: the automatic choice splits K over two pairs
    try:
        for M, N, K, dt in cases:
            torch.manual_seed(M + N + K + mode)
            x = torch.randn(M, K, device="cuda", dtype=dt)
            w = torch.randn(N, K, device="cuda", dtype=dt) / math.sqrt(K)
            b = torch.randn(N, device="cuda", dtype=dt)
            res = torch.randn(M, N, device="cuda", dtype=dt)
            flavours = [dict(), dict(bias=b, act=R.ACT_GELU_TANH, act_start=(N // 256) * 128), dict(bias=b, residual=res),
                        dict(act=R.ACT_SILU, gated=True), dict(out_dtype=torch.float32)]
            for kw in flavours:
                lib.llmss_gemm_set_pair(mode)
                y = cu.linear(x, w, **kw)
                y_again = cu.linear(x, w, **kw)
                lib.llmss_gemm_set_pair(0)
                y_single = cu.linear(x, w, **kw)
                torch.cuda.synchronize()
                ref = R.linear(x, w, **{**kw, "out_dtype": torch.float32})
                tag = f"pair mode {mode} {M}x{N}x{K} {sorted(kw)}"
                _close(y, ref, 4e-2 if dt == torch.bfloat16 else 8e-3, 2e-2, tag)
                assert torch.equal(y, y_again), tag + ": not repeatable"
                assert (y.float() - y_single.float()).abs().max().item() <= 0.05 * max(1.0, ref.abs().max().item()), tag
    finally:
        lib.llmss_gemm_set_pair(-1)


# ------------------------------------------------------------------ norm
@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("kind", [R.NORM_LAYER, R.NORM_RMS])
@pytest.mark.parametrize("H", [768, 4096, 6144])
def test_norm_residual(H, kind, dtype):
    cu = _cu()
    M = 37
    x = torch.randn(M, H, device="cuda", dtype=dtype)
    res = torch.randn(M, H, device="cuda", dtype=dtype)
    x2 = torch.randn(M, H, device="cuda", dtype=dtype)
    w = torch.randn(H, device="cuda", dtype=dtype)
    b = torch.randn(H, device="cuda", dtype=dtype) if kind == R.NORM_LAYER else None
    y, r = cu.norm_residual(x, res, w, b, 1e-5, kind, x2=x2)
    s = (x.float() + x2.float() + res.float())
    _close(r, s, 4e-2, 1e-2, "residual")
    yr, _ = R.norm_residual(r, None, w, b, 1e-5, kind, out_dtype=torch.float32)
    _close(y, yr, 5e-2, 2e-2, "normed")
    y0, r0 = cu.norm_residual(x, None, w, b, 1e-5, kind)
    assert r0 is x
    _close(y0, R.norm_residual(x, None, w, b, 1e-5, kind, out_dtype=torch.float32)[0], 5e-2, 2e-2, "no residual")


# ------------------------------------------------------------------ rope + kv write
@pytest.mark.parametrize("style,d,rd", [(R.ROPE_GPTJ, 256, 64), (R.ROPE_NEOX, 128, 128), (R.ROPE_NONE, 64, 0),
                                        (R.ROPE_GPTJ, 64, 64)])
def test_rope_kv_write(style, d, rd):
    cu = _cu()
    T, hq, hkv, bs, nb = 19, 4, 2, 16, 12
    W = (hq + 2 * hkv) * d + 64
    qkv = torch.randn(T, W, device="cuda", dtype=torch.bfloat16)
    pos = torch.randint(0, 3000, (T,), device="cuda", dtype=torch.int32)
    slots = torch.randperm(nb * bs, device="cuda")[:T].to(torch.int32)
    slots[3] = -1
    kc = torch.zeros(nb, hkv, bs, d, device="cuda", dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    q_off, k_off, v_off = 64, 64 + hq * d, 64 + (hq + hkv) * d
    ref_qkv, rk, rv = qkv.clone(), kc.clone(), vc.clone()
    R.rope_kv_write(ref_qkv, pos, slots, rk, rv, hq, hkv, d, rd, style, 10000.0, q_off, k_off, v_off)
    cu.rope_kv_write(qkv, pos, slots, kc, vc, hq, hkv, d, rd, style, 10000.0, q_off, k_off, v_off)
    _close(qkv, ref_qkv, 3e-2, 2e-2, "qkv")
    _close(kc, rk, 3e-2, 2e-2, "k cache")
    _close(vc, rv, 0, 0, "v cache")


# ------------------------------------------------------------------ decode attention
@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("hq,hkv,d", [(4, 4, 256), (8, 8, 128), (12, 12, 64), (8, 2, 128), (6, 1, 128), (16, 1, 64),
                                      (48, 1, 128), (32, 8, 128), (16, 4, 256), (6, 2, 64)])
@pytest.mark.parametrize("splits", [0, 1, 4])  # 0 = balanced flat partition (tensor-core kernel only)
def test_attn_decode(hq, hkv, d, splits, dtype):
    cu = _cu()
    torch.manual_seed(0)
    B, bs = 5, 16
    lens = [1, 17, 300, 64, 1025]
    maxb = (max(lens) + bs - 1) // bs
    nb = 1 + B * maxb
    kc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=dtype)
    vc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=dtype)
    perm = (torch.randperm(nb - 1) + 1).tolist()
    tables = [[perm[i * maxb + j] for j in range(maxb)] for i in range(B)]
    bt = torch.tensor(tables, dtype=torch.int32, device="cuda")
    sl = torch.tensor(lens, dtype=torch.int32, device="cuda")
    W = hq * d + 128
    qbuf = torch.randn(B, W, device="cuda", dtype=dtype)
    out = torch.zeros(B, W, device="cuda", dtype=dtype)
    q_view, o_view = qbuf[:, 128:], out[:, 128:]
    cu.attn_decode(q_view, kc, vc, bt, sl, o_view, hq, hkv, d, d ** -0.5, splits)
    cu_q = torch.arange(B + 1, dtype=torch.int32, device="cuda")
    ref = R.attention_paged(q_view.reshape(B, hq, d), kc, vc, bt, sl, cu_q, d ** -0.5).reshape(B, hq * d)
    _close(o_view, ref, 2e-2, 2e-2, "attn_decode")
    assert out[:, :128].abs().sum().item() == 0


@pytest.mark.parametrize("bs", [16, 32, 64])
def test_attn_decode_tc_matches_cuda_core_path(monkeypatch, bs):
    """Grouped heads: the tcgen05 kernel and the CUDA-core kernel must agree (and both be reachable)."""
    cu = _cu()
    torch.manual_seed(1)
    B, hq, hkv, d = 7, 16, 2, 128
    lens = [5, 128, 129, 640, 1, 2047, 256]
    maxb = (max(lens) + bs - 1) // bs
    nb = 1 + B * maxb
    kc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=torch.bfloat16)
    vc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=torch.bfloat16)
    perm = (torch.randperm(nb - 1) + 1).tolist()
    bt = torch.tensor([[perm[i * maxb + j] for j in range(maxb)] for i in range(B)], dtype=torch.int32, device="cuda")
    sl = torch.tensor(lens, dtype=torch.int32, device="cuda")
    q = torch.randn(B, hq * d, device="cuda", dtype=torch.bfloat16)
    outs = []
    for min_group in (4, 1000):  # tensor-core path, then CUDA-core path
        monkeypatch.setattr(cu, "_TC_MIN_GROUP", min_group)
        assert cu.use_tc_decode(hq, hkv, d) == (min_group == 4)
        for splits in (0, 1, 3, 8):
            o = torch.zeros_like(q)
            cu.attn_decode(q, kc, vc, bt, sl, o, hq, hkv, d, d ** -0.5, splits)
            outs.append(o.float())
    cu_q = torch.arange(B + 1, dtype=torch.int32, device="cuda")
    ref = R.attention_paged(q.reshape(B, hq, d), kc, vc, bt, sl, cu_q, d ** -0.5).reshape(B, hq * d)
    for o in outs:
        _close(o, ref, 2e-2, 2e-2, "attn_decode tc / cuda-core vs oracle")


@pytest.mark.parametrize("ws", ["1", "0"])
@pytest.mark.parametrize("hq,hkv,d", [(16, 16, 256), (8, 8, 128), (8, 2, 128)])
def test_attn_decode_ws_and_plain_mma(hq, hkv, d, ws, monkeypatch):
    """Groups of <= 8 q heads on 64-token pages: the weight-stationary 32-row MMA form (default) and the plain
    64-row form (LLMSS_ATTN_WS=0) of the tcgen05 decode kernel must both match the oracle."""
    cu = _cu()
    monkeypatch.setenv("LLMSS_ATTN_WS", ws)
    torch.manual_seed(2)
    B, bs = 6, 64
    lens = [1, 64, 65, 300, 1025, 130]
    maxb = (max(lens) + bs - 1) // bs
    nb = 1 + B * maxb
    kc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=torch.bfloat16)
    vc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=torch.bfloat16)
    perm = (torch.randperm(nb - 1) + 1).tolist()
    bt = torch.tensor([[perm[i * maxb + j] for j in range(maxb)] for i in range(B)], dtype=torch.int32, device="cuda")
    sl = torch.tensor(lens, dtype=torch.int32, device="cuda")
    q = torch.randn(B, hq * d, device="cuda", dtype=torch.bfloat16)
    cu_q = torch.arange(B + 1, dtype=torch.int32, device="cuda")
    ref = R.attention_paged(q.reshape(B, hq, d), kc, vc, bt, sl, cu_q, d ** -0.5).reshape(B, hq * d)
    for splits in (1, 3):
        o = torch.zeros_like(q)
        cu.attn_decode(q, kc, vc, bt, sl, o, hq, hkv, d, d ** -0.5, splits)
        torch.cuda.synchronize()
        _close(o, ref, 2e-2, 2e-2, f"attn_decode ws={ws} splits={splits}")


# ------------------------------------------------------------------ embed / sample
def test_embed():
    cu = _cu()
    V, H, P, T = 1000, 768, 128, 33
    wte = torch.randn(V, H, device="cuda", dtype=torch.bfloat16)
    wpe = torch.randn(P, H, device="cuda", dtype=torch.bfloat16)
    tok = torch.randint(0, V, (T,), device="cuda", dtype=torch.int32)
    pos = torch.randint(0, P - 2, (T,), device="cuda", dtype=torch.int32)
    _close(cu.embed(tok, wte), R.embed(tok, wte), 0, 0, "wte")
    _close(cu.embed(tok, wte, pos, wpe, 2), R.embed(tok, wte, pos, wpe, 2).float(), 2e-2, 1e-2, "wte+wpe")


def test_vocab_parallel_argmax_kernels():
    """K11 greedy case on one GPU: per-shard candidates + merge == argmax of the full logits (first maximum wins)."""
    import ctypes as C

    from llmss_b200.ops import _lib

    lib = _lib.load()
    torch.manual_seed(5)
    B, V, t = 37, 50400, 8
    Vb = -(-V // t)  # ceil-div shards, the last one is partly padding
    logits = torch.randn(B, Vb * t, device="cuda")
    logits[:, V:] = 1e9  # padding columns must never win
    logits[3, 100] = logits[3, 40000] = 50.0   # tie across shards -> lowest index
    logits[4, 7] = logits[4, 8] = 60.0         # tie inside a shard
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    cands = torch.empty(B, t, 4, dtype=torch.int32, device="cuda")
    for r in range(t):
        shard = logits[:, r * Vb:(r + 1) * Vb]
        c = torch.empty(B, 4, dtype=torch.int32, device="cuda")
        n_valid = max(0, min(Vb, V - r * Vb))
        _lib.check(lib.llmss_argmax_local(C.c_void_p(shard.data_ptr()), shard.stride(0), B, n_valid, r * Vb,
                                          C.c_void_p(c.data_ptr()), stream), "argmax_local")
        cands[:, r] = c
    out = torch.empty(B, dtype=torch.int32, device="cuda")
    _lib.check(lib.llmss_argmax_merge(C.c_void_p(cands.data_ptr()), t, B, C.c_void_p(out.data_ptr()), None, stream),
               "argmax_merge")
    ref = logits[:, :V].argmax(dim=-1)
    assert out.tolist()[3] == 100 and out.tolist()[4] == 7
    rows = [i for i in range(B) if i not in (3, 4)]
    assert torch.equal(out.long()[rows], ref[rows])


def _oracle_choices(fn, rng):
    """The oracle's token per row, plus the tokens it would pick with the uniform draw nudged by +-2e-4: the CUDA
    kernels evaluate the same CDF in fp32 (`__expf`), so a row whose draw sits on a CDF boundary may legitimately
    land on the neighbouring token; every other row must match exactly."""
    want = fn(rng.clone(), 0.0).long().cpu()
    lo, hi = fn(rng.clone(), -2e-4).long().cpu(), fn(rng.clone(), 2e-4).long().cpu()
    return want, lo, hi


def _check_tokens(got, want, lo, hi, what):
    got = got.long().cpu()
    exact = got == want
    near = exact | (got == lo) | (got == hi)
    assert bool(near.all()), f"{what}: rows {(~near).nonzero().flatten().tolist()} differ from the oracle\n{got}\n{want}"
    assert exact.float().mean().item() >= 0.9, f"{what}: only {exact.float().mean().item():.2f} of rows match exactly"


def test_sample_matches_oracle():
    cu = _cu()
    torch.manual_seed(3)
    B, V = 16, 50400
    logits = (torch.randn(B, V + 32, device="cuda") * 3).float()
    greedy = torch.tensor([1, 0] * (B // 2), dtype=torch.int32, device="cuda")
    temp = torch.tensor([1.0, 0.7, 1.0, 0.5] * (B // 4), device="cuda")
    top_k = torch.tensor([50, 50, 0, 5, 0, 1000, 1, 0] * (B // 8), dtype=torch.int32, device="cuda")
    top_p = torch.tensor([0.95, 0.95, 1.0, 0.5, 0.9, 1.0, 1.0, 1.0] * (B // 8), device="cuda")
    for step in range(6):
        rng = torch.tensor([[1234 + 17 * b, step] for b in range(B)], dtype=torch.int32, device="cuda")
        want, lo, hi = _oracle_choices(lambda r, sh: R.sample(logits, greedy, temp, top_k, top_p, r, vocab=V, u_shift=sh), rng)
        rng_ref = rng.clone()
        R.sample(logits, greedy, temp, top_k, top_p, rng_ref, vocab=V)
        got = cu.sample(logits, greedy, temp, top_k, top_p, rng, vocab=V)
        torch.cuda.synchronize()
        assert torch.equal(rng, rng_ref.to(rng.dtype))
        _check_tokens(got, want, lo, hi, f"step {step}")
        g = greedy.bool()
        assert torch.equal(got.long()[g], logits[g, :V].argmax(-1))
        assert int(got.max()) < V


def test_sample_plain_multinomial_never_falls_through():
    """top_k = 0, top_p = 1: every thread must see the same draw (the RNG counter is advanced only after all
    threads have read it) and the per-thread CDF intervals must tile [0, total) without gaps -- before the fix a
    missed interval returned token vocab - 1."""
    cu = _cu()
    torch.manual_seed(5)
    B, V = 64, 50257
    logits = (torch.randn(B, V, device="cuda") * 4).float()
    logits[:, V - 1] = -60.0   # the old fall-through token: probability ~ 0
    z = torch.zeros(B, dtype=torch.int32, device="cuda")
    one = torch.ones(B, device="cuda")
    for step in range(50):
        rng = torch.tensor([[77 + b, step] for b in range(B)], dtype=torch.int32, device="cuda")
        rng0 = rng.clone()   # the kernel advances the counter in place
        got = cu.sample(logits, z, one, z, one, rng, vocab=V)
        assert int((got == V - 1).sum()) == 0, f"step {step}: fell through to the last token"
        rng = rng0
        if step < 4:
            # index-order CDF over 50k tokens, most of them with weights near the fp32 rounding of the running sum:
            # the pick may move by a few tokens, but it must stay between the oracle's picks for u -+ 2e-4
            _, lo, hi = _oracle_choices(lambda r, sh: R.sample(logits, z, one, z, one, r, vocab=V, u_shift=sh), rng)
            g = got.long().cpu()
            assert bool(((g >= lo) & (g <= hi)).all()), f"multinomial step {step}: picks outside the oracle's bracket\n{g}\n{lo}\n{hi}"


@pytest.mark.parametrize("t", [2, 4, 8])
def test_vocab_parallel_sampling_kernels(t):
    """K11 general case on one GPU: the t ranks' shard summaries are produced one after the other, concatenated the
    way the record all-gather lays them out, and merged; tokens must equal the vocab-parallel oracle (and therefore
    the single-GPU sampler on the gathered logits)."""
    cu = _cu()
    torch.manual_seed(11 + t)
    B, V = 24, 50257                      # V not divisible by t: the last shard is padded
    Vb = -(-V // t)
    full = (torch.randn(B, Vb * t, device="cuda") * 3).float()
    full[:, V:] = 1e9                     # padding columns hold garbage that must never win
    full[3, 100] = full[3, 40000] = full[3, :V].max() + 1.0   # exact tie across two shards (greedy row): lowest index wins
    greedy = torch.tensor([1, 0, 0] * (B // 3), dtype=torch.int32, device="cuda")
    temp = torch.tensor([1.0, 0.7, 0.5, 1.0] * (B // 4), device="cuda")
    top_k = torch.tensor([50, 5, 128, 1, 64, 20] * (B // 6), dtype=torch.int32, device="cuda")
    top_p = torch.tensor([0.95, 0.5, 1.0, 0.9] * (B // 4), device="cuda")
    plain_rows = [1, 8, 13, 22]            # drawn from the whole distribution: top_k = 0, top_p = 1
    for r_ in plain_rows:
        top_k[r_], top_p[r_], greedy[r_] = 0, 1.0, 0
    shards = [full[:, r * Vb:(r + 1) * Vb].contiguous() for r in range(t)]
    cpu = lambda x: x.cpu()  # noqa: E731
    for step in range(4):
        rng = torch.tensor([[900 + 13 * b, step] for b in range(B)], dtype=torch.int32, device="cuda")
        want, lo, hi = _oracle_choices(
            lambda r, sh: R.sample_vocab_parallel([cpu(x) for x in shards], cpu(greedy), cpu(temp), cpu(top_k), cpu(top_p),
                                                  cpu(r), V, k_summary=cu.VP_KS, u_shift=sh), rng)
        nv = [max(0, min(Vb, V - r * Vb)) for r in range(t)]
        recs = [cu.vp_summary(shards[r], nv[r], r * Vb, greedy, top_k, temperature=temp) for r in range(t)]
        allr = torch.cat(recs, dim=1).contiguous()
        cands = [cu.vp_multinomial_pick(shards[r], nv[r], r * Vb, allr, t, r, greedy, temp, top_k, top_p, rng) for r in range(t)]
        got = cu.vp_merge_sample(allr, t, V, greedy, temp, top_k, top_p, rng)
        cu.vp_multinomial_merge(torch.cat(cands, dim=1).contiguous(), t, got)
        torch.cuda.synchronize()
        assert rng[:, 1].tolist() == [step + 1] * B
        call = torch.stack([c[:, 0] for c in cands], 1).cpu()      # exactly one rank names a token for a plain row
        for b in range(B):
            n_named = int((call[b] >= 0).sum())
            assert n_named == (1 if b in plain_rows else 0), (b, call[b].tolist())
        other = [b for b in range(B) if b not in plain_rows]
        _check_tokens(got[other], want[other], lo[other], hi[other], f"t={t} step {step}")
        g = got.long().cpu()                                       # index-order CDF: bracket instead of equality
        for b in plain_rows:
            assert min(lo[b], hi[b]) <= g[b] <= max(lo[b], hi[b]), (t, step, b, int(g[b]), int(lo[b]), int(hi[b]))
        assert int(got.max()) < V
        single = cu.sample(full[:, :V].contiguous(), greedy, temp, top_k, top_p,
                           torch.tensor([[900 + 13 * b, step] for b in range(B)], dtype=torch.int32, device="cuda"), vocab=V)
        assert (single.long()[other] == got.long()[other]).float().mean().item() >= 0.9
    assert got.tolist()[3] == 100


@pytest.mark.parametrize("mode", ["streamk", "cluster"])
def test_gemm_modes_agree(mode):
    """Both schedulers (persistent stream-K, thread-block-cluster split-K) in a fresh process."""
    import os
    import subprocess
    import sys

    code = r"""
import math, torch
from llmss_b200.ops import cuda_impl as cu, reference as R
torch.manual_seed(0)
bad = 0
for (M, N, K) in [(1, 128, 128), (8, 4096, 2560), (16, 3584, 4096), (64, 4096, 2560), (64, 4096, 20480), (33, 1000, 328),
                  (128, 2048, 1024), (256, 4096, 4096), (64, 512, 8192)]:
    for dt in (torch.bfloat16, torch.float16):
        x = torch.randn(M, K, device='cuda', dtype=dt); w = torch.randn(N, K, device='cuda', dtype=dt) / math.sqrt(K)
        b = torch.randn(N, device='cuda', dtype=dt); r = torch.randn(M, N, device='cuda', dtype=dt)
        p = cu.gemm_plan(M, N, K)
        y = cu.linear(x, w, b, act=R.ACT_GELU_TANH, residual=r)
        y2 = cu.linear(x, w, b, act=R.ACT_GELU_TANH, residual=r)
        ref = R.linear(x, w, b, act=R.ACT_GELU_TANH, residual=r, out_dtype=torch.float32)
        err = (y.float() - ref).abs()
        nbad = int((err > 4e-2 + 2e-2 * ref.abs()).sum())
        same = torch.equal(y, y2)
        print(M, N, K, dt, 'split', p.split, 'grid', p.grid, 'max err', err.max().item(), 'bad', nbad, 'repeatable', same)
        bad += nbad + (0 if same else 1)
    g = cu.linear(x[:, :], torch.randn(N + (N % 2), K, device='cuda', dtype=dt) / math.sqrt(K), act=R.ACT_SILU, gated=True)
torch.cuda.synchronize()
print('MODES_OK' if bad == 0 else 'MODES_FAILED')
"""
    env = dict(os.environ, LLMSS_GEMM_MODE=mode, PYTHONPATH=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, env=env)
    assert "MODES_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-2000:]


# ------------------------------------------------------------------ prefill attention (tcgen05)
@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("hq,hkv,d", [(4, 4, 256), (8, 8, 128), (6, 6, 64), (8, 2, 128), (6, 1, 128)])
def test_attn_prefill(hq, hkv, d, dtype):
    cu = _cu()
    torch.manual_seed(1)
    lens = [1, 5, 128, 129, 300, 64, 700]
    T = sum(lens)
    pad = 64
    W = pad + (hq + 2 * hkv) * d
    qkv = torch.randn(T, W, device="cuda", dtype=dtype)
    q_off, k_off, v_off = pad, pad + hq * d, pad + (hq + hkv) * d
    out = torch.zeros(T, hq * d + 32, device="cuda", dtype=dtype)
    cu_seq = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device="cuda")
    scale = d ** -0.5
    cu.attn_prefill(qkv[:, q_off:q_off + hq * d], qkv[:, k_off:k_off + hkv * d], qkv[:, v_off:v_off + hkv * d],
                    out[:, :hq * d], cu_seq, len(lens), max(lens), hq, hkv, d, scale)
    torch.cuda.synchronize()
    g = hq // hkv
    ref = torch.zeros(T, hq * d, device="cuda", dtype=torch.float32)
    s0 = 0
    for L in lens:
        q = qkv[s0:s0 + L, q_off:q_off + hq * d].float().reshape(L, hq, d).permute(1, 0, 2)
        k = qkv[s0:s0 + L, k_off:k_off + hkv * d].float().reshape(L, hkv, d).permute(1, 0, 2).repeat_interleave(g, 0)
        v = qkv[s0:s0 + L, v_off:v_off + hkv * d].float().reshape(L, hkv, d).permute(1, 0, 2).repeat_interleave(g, 0)
        s = (q @ k.transpose(1, 2)) * scale
        s = s.masked_fill(torch.triu(torch.ones(L, L, dtype=torch.bool, device="cuda"), 1), float("-inf"))
        ref[s0:s0 + L] = (torch.softmax(s, -1) @ v).permute(1, 0, 2).reshape(L, hq * d)
        s0 += L
    _close(out[:, :hq * d], ref, 3e-2, 3e-2, "attn_prefill")
    assert out[:, hq * d:].abs().sum().item() == 0


@pytest.mark.parametrize("bs", [64, 16])
@pytest.mark.parametrize("hq,hkv,d", [(4, 4, 256), (8, 2, 128), (6, 6, 64), (6, 1, 128)])
def test_attn_prefill_over_context(hq, hkv, d, bs):
    """Multi-token queries over prior context (chunked prefill / prefix reuse): K / V come from the paged cache
    through the block table; compared with the fp32 gather oracle.  Mixes pure prefill (ctx 0), a decode-like
    row (1 query), chunks that start and end inside pages, and a chunk longer than one query tile."""
    cu = _cu()
    torch.manual_seed(4)
    dtype = torch.bfloat16
    ctx = [0, 37, 64, 500, 129, 1000, 5]
    new = [70, 1, 128, 300, 63, 17, 200]
    B = len(ctx)
    tot = [c + n for c, n in zip(ctx, new)]
    maxb = max(-(-t // bs) for t in tot) + 1
    nb = 1 + B * maxb
    kc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=dtype)
    vc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=dtype)
    perm = (torch.randperm(nb - 1) + 1).tolist()
    bt = torch.zeros(B, maxb, dtype=torch.int32)
    for i in range(B):
        n = -(-tot[i] // bs)
        bt[i, :n] = torch.tensor(perm[i * maxb:i * maxb + n], dtype=torch.int32)
    bt = bt.cuda()
    sl = torch.tensor(tot, dtype=torch.int32, device="cuda")
    cu_q = torch.tensor([0] + list(torch.tensor(new).cumsum(0)), dtype=torch.int32, device="cuda")
    T = sum(new)
    W = 32 + hq * d
    qbuf = torch.randn(T, W, device="cuda", dtype=dtype)
    q = qbuf[:, 32:]
    out = torch.zeros(T, hq * d + 16, device="cuda", dtype=dtype)
    scale = d ** -0.5
    cu.attn_prefill_paged(q, kc, vc, bt, sl, out[:, :hq * d], cu_q, B, max(new), hq, hkv, d, scale)
    torch.cuda.synchronize()
    ref = R.attention_paged(q.float().reshape(T, hq, d), kc.float(), vc.float(), bt, sl, cu_q, scale).reshape(T, hq * d)
    _close(out[:, :hq * d], ref, 3e-2, 3e-2, "attn_prefill_paged")
    assert out[:, hq * d:].abs().sum().item() == 0
    # in place over q, as the decoder calls it
    cu.attn_prefill_paged(q, kc, vc, bt, sl, q, cu_q, B, max(new), hq, hkv, d, scale)
    torch.cuda.synchronize()
    assert torch.equal(q, out[:, :hq * d])


def test_attn_prefill_in_place_over_q():
    """The fused GPT-J block lets the context overwrite q inside the packed activation."""
    cu = _cu()
    torch.manual_seed(2)
    hq, hkv, d, lens = 4, 4, 256, [200, 333]
    T = sum(lens)
    qkv = torch.randn(T, (hq + 2 * hkv) * d, device="cuda", dtype=torch.bfloat16)
    k_off, v_off, q_off = 0, hkv * d, 2 * hkv * d
    ref = torch.empty(T, hq * d, device="cuda", dtype=torch.bfloat16)
    cu_seq = torch.tensor([0, 200, 533], dtype=torch.int32, device="cuda")
    args = (cu_seq, 2, 333, hq, hkv, d, d ** -0.5)
    cu.attn_prefill(qkv[:, q_off:], qkv[:, k_off:k_off + hkv * d], qkv[:, v_off:v_off + hkv * d], ref, *args)
    cu.attn_prefill(qkv[:, q_off:], qkv[:, k_off:k_off + hkv * d], qkv[:, v_off:v_off + hkv * d], qkv[:, q_off:], *args)
    torch.cuda.synchronize()
    assert torch.equal(qkv[:, q_off:], ref)
