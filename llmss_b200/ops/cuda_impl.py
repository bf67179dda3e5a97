"""CUDA execution path: thin wrappers that hand torch tensors to the sm_100a kernels.

All launches use torch's current stream, so they are captured by ``torch.cuda.graph``.
Outputs are allocated with torch (graph-pool friendly); scratch for stream-K / split-KV is
a per-device arena allocated once.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from . import _lib
from .reference import ACT_NONE

_DT = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2}


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _count(n: int = 1):
    _lib.LAUNCHES["count"] += n


class _Arena:
    """Per-device scratch: stream-K partials + tile counters + split-KV partials."""

    SK_FLOATS = 2 * 148 * 128 * 256
    SK_COUNTERS = 1 << 16

    def __init__(self, device):
        self.sk_ws = torch.empty(self.SK_FLOATS, dtype=torch.float32, device=device)
        self.sk_counters = torch.zeros(self.SK_COUNTERS, dtype=torch.int32, device=device)
        self.attn = None
        self.ag_maps = None
        self._retired = []  # outgrown buffers stay alive: captured CUDA graphs may still point at them

    def attn_scratch(self, nfloats: int, device):
        if self.attn is None or self.attn.numel() < nfloats:
            if self.attn is not None:
                self._retired.append(self.attn)
            self.attn = torch.empty(max(nfloats, 1 << 22), dtype=torch.float32, device=device)
        return self.attn


_ARENAS = {}


def arena(device) -> _Arena:
    key = torch.device(device).index or 0
    a = _ARENAS.get(key)
    if a is None:
        a = _ARENAS[key] = _Arena(device)
    return a


def _rows2d(x: torch.Tensor):
    assert x.dim() == 2 and x.stride(1) == 1, "expected a 2-D tensor with contiguous last dim"
    return x.shape[0], x.shape[1], x.stride(0)


# --------------------------------------------------------------------------- GEMM
def gemm_plan(M: int, N: int, K: int, nranks: int = 1, max_ctas: int = 0) -> _lib.GemmPlan:
    p = _lib.GemmPlan()
    _lib.check(_lib.load().llmss_gemm_plan_query(M, N, K, nranks, 2, max_ctas, C.byref(p)), "gemm_plan_query")
    return p


def linear(x, w, bias=None, act: int = ACT_NONE, gated: bool = False, act_start: int = 0, residual=None,
           out_dtype=None, symm=None, out=None, max_ctas: int = 0, collective: int = 1):
    """``symm``: fuse the tensor-parallel all-reduce into the epilogue.  ``collective``: 1 = library picks the
    exchange (in-switch NVLS reduction from `LLMSS_NVLS_MIN_BYTES` upwards when the workspace is multicast
    bound, else the one-shot Lamport push), 4 = NVLS, 5 = one-shot push."""
    lib = _lib.load()
    M, K, ldx = _rows2d(x)
    N, Kw, ldw = _rows2d(w)
    assert K == Kw, f"linear: x has K={K}, w has K={Kw}"
    assert x.dtype == w.dtype and x.dtype in (torch.float16, torch.bfloat16)
    out_f32 = 1 if out_dtype == torch.float32 else 0
    n_out = N // 2 if gated else N
    if out is None:
        out = torch.empty(M, n_out, dtype=torch.float32 if out_f32 else x.dtype, device=x.device)
    if M == 0:
        return out
    ldr = residual.stride(0) if residual is not None else 0
    ar = symm is not None and symm.nranks > 1
    a = arena(x.device)
    rc = lib.llmss_gemm_ex(_ptr(x), ldx, _ptr(w), ldw, _ptr(bias), _ptr(residual), ldr, _ptr(out), out.stride(0),
                           M, N, K, _DT[x.dtype], out_f32, act, act_start, 1 if gated else 0,
                           _ptr(a.sk_ws), _ptr(a.sk_counters), symm.handle if ar else None, max_ctas,
                           collective if ar else 0, None, _stream())
    _lib.check(rc, "llmss_gemm")
    _count()
    return out


def linear_reduce_scatter(x, w, symm, bias=None, residual_shard=None):
    """K2: row-parallel GEMM whose epilogue reduce-scatters over token rows.  Returns this rank's
    [M / t, N] shard of sum_r(x_r @ w_r.T) (+ bias, + residual_shard)."""
    lib = _lib.load()
    M, K, ldx = _rows2d(x)
    N = w.shape[0]
    rows = M // symm.nranks
    out = torch.empty(rows, N, dtype=x.dtype, device=x.device)
    a = arena(x.device)
    rc = lib.llmss_gemm_ex(_ptr(x), ldx, _ptr(w), w.stride(0), _ptr(bias), _ptr(residual_shard),
                           residual_shard.stride(0) if residual_shard is not None else 0, _ptr(out), out.stride(0),
                           M, N, K, _DT[x.dtype], 0, ACT_NONE, 0, 0, _ptr(a.sk_ws), _ptr(a.sk_counters), symm.handle, 0,
                           2, None, _stream())
    _lib.check(rc, "llmss_gemm_ex(reduce-scatter)")
    _count()
    return out


def allgather_linear(x_shard, w, symm, bias=None, act: int = ACT_NONE, gated: bool = False, act_start: int = 0,
                     out_dtype=None):
    """K3: column-parallel GEMM over the all-gather of the ranks' row shards; the gather is fused into
    the kernel's TMA loads (token tiles are pulled from the owning rank's HBM over NVLink)."""
    lib = _lib.load()
    rows, K, ldx = _rows2d(x_shard)
    M, N = rows * symm.nranks, w.shape[0]
    out_f32 = 1 if out_dtype == torch.float32 else 0
    out = torch.empty(M, N // 2 if gated else N, dtype=torch.float32 if out_f32 else x_shard.dtype, device=x_shard.device)
    a = arena(x_shard.device)
    if a.ag_maps is None:
        a.ag_maps = torch.zeros(2 * 8 * 128 + 64, dtype=torch.uint8, device=x_shard.device)
    maps_ptr = (a.ag_maps.data_ptr() + 63) // 64 * 64
    rc = lib.llmss_gemm_ex(_ptr(x_shard), ldx, _ptr(w), w.stride(0), _ptr(bias), None, 0, _ptr(out), out.stride(0),
                           M, N, K, _DT[x_shard.dtype], out_f32, act, act_start, 1 if gated else 0, _ptr(a.sk_ws),
                           _ptr(a.sk_counters), symm.handle, 0, 3, C.c_void_p(maps_ptr), _stream())
    _lib.check(rc, "llmss_gemm_ex(all-gather)")
    _count()
    return out


# --------------------------------------------------------------------------- norm
def norm_residual(x, residual, weight, bias, eps: float, kind: int, x2=None, write_residual: bool = True):
    lib = _lib.load()
    M, H, ldx = _rows2d(x)
    out = torch.empty(M, H, dtype=x.dtype, device=x.device)
    need_res = write_residual and (residual is not None or x2 is not None)
    new_res = torch.empty(M, H, dtype=x.dtype, device=x.device) if need_res else None
    # the kernel uses one stride for residual in and out
    if residual is not None and new_res is not None and residual.stride(0) != new_res.stride(0):
        residual = residual.contiguous()
    ldr = new_res.stride(0) if new_res is not None else (residual.stride(0) if residual is not None else H)
    rc = lib.llmss_norm_residual(_ptr(x), _ptr(x2), _ptr(residual), _ptr(new_res), _ptr(weight), _ptr(bias), _ptr(out),
                                 M, H, ldx, x2.stride(0) if x2 is not None else H, ldr, out.stride(0), eps, kind,
                                 _DT[x.dtype], _stream())
    _lib.check(rc, "llmss_norm_residual")
    _count()
    return out, (new_res if new_res is not None else x)


# --------------------------------------------------------------------------- rope + kv
def rope_kv_write(qkv, positions, slot_mapping, k_cache, v_cache, hq, hkv, d, rotary_dim, style, base,
                  q_off, k_off, v_off):
    lib = _lib.load()
    T, _, ld = _rows2d(qkv)
    bs = k_cache.shape[2]
    rc = lib.llmss_rope_kv_write(_ptr(qkv), _ptr(positions), _ptr(slot_mapping), _ptr(k_cache), _ptr(v_cache), T, ld,
                                 hq, hkv, d, rotary_dim, style, float(base), q_off, k_off, v_off, bs, _DT[qkv.dtype],
                                 _stream())
    _lib.check(rc, "llmss_rope_kv_write")
    _count()


# --------------------------------------------------------------------------- attention
def decode_splits(B: int, hq: int, hkv: int, d: int, max_kv: int, dtype=torch.bfloat16, block_size: int = 16) -> int:
    """Split-KV factor chosen by the library so the grid fills whole waves of resident CTAs."""
    if d not in (64, 128, 256):
        return 1
    if use_tc_decode(hq, hkv, d, block_size):
        return int(_lib.load().llmss_attn_decode_tc_pick_splits(B, hkv, d, max_kv))
    return int(_lib.load().llmss_attn_decode_pick_splits(B, hq, hkv, d, max_kv, _DT[dtype]))


_TC_MIN_GROUP = int(os.environ.get("LLMSS_ATTN_TC_MIN_GROUP", "4"))


def use_tc_decode(hq: int, hkv: int, d: int, block_size: int = 16) -> bool:
    """tcgen05 decode attention (q heads of a KV group = MMA rows): always with 64-token pages (one TMA op
    per tile: 95 % of HBM peak even for MHA), and for grouped / multi-query models with smaller pages."""
    g = hq // max(hkv, 1)
    if g > 128 or d not in (64, 128, 256) or block_size not in (16, 32, 64):
        return False
    return block_size == 64 or g >= _TC_MIN_GROUP


def attn_decode(q_view, k_cache, v_cache, block_tables, seq_lens, out_view, hq, hkv, d, scale, num_splits):
    lib = _lib.load()
    B = q_view.shape[0]
    bs = k_cache.shape[2]
    scratch = None
    tc = use_tc_decode(hq, hkv, d, bs)
    if num_splits == 0 and not tc:
        num_splits = 1  # balanced partition exists only in the tensor-core kernel
    if num_splits > 1:
        n = lib.llmss_attn_decode_scratch_floats(B, hq, d, num_splits)
        scratch = arena(q_view.device).attn_scratch(n, q_view.device)
    elif num_splits == 0:
        n = lib.llmss_attn_decode_tc_balanced_scratch_floats(B, hq, d, block_tables.shape[1], bs)
        scratch = arena(q_view.device).attn_scratch(n, q_view.device)
    if tc:
        rc = lib.llmss_attn_decode_tc(_ptr(q_view), q_view.stride(0), _ptr(k_cache), _ptr(v_cache), _ptr(block_tables),
                                      block_tables.shape[1], _ptr(seq_lens), _ptr(out_view), out_view.stride(0),
                                      _ptr(scratch), B, hq, hkv, d, bs, k_cache.shape[0], float(scale), num_splits,
                                      _DT[q_view.dtype], _stream())
        _lib.check(rc, "llmss_attn_decode_tc")
        _count(1 if num_splits == 1 else 2)
        return
    rc = lib.llmss_attn_decode(_ptr(q_view), q_view.stride(0), _ptr(k_cache), _ptr(v_cache), _ptr(block_tables),
                               block_tables.shape[1], _ptr(seq_lens), _ptr(out_view), out_view.stride(0),
                               _ptr(scratch), B, hq, hkv, d, bs, float(scale), num_splits, _DT[q_view.dtype],
                               _stream())
    _lib.check(rc, "llmss_attn_decode")
    _count(2 if num_splits > 1 else 1)


def attn_prefill(q, k, v, out, cu_seqlens, num_seqs, max_seqlen, hq, hkv, d, scale):
    lib = _lib.load()
    rc = lib.llmss_attn_prefill(_ptr(q), q.stride(0), _ptr(k), k.stride(0), _ptr(v), v.stride(0), _ptr(out),
                                out.stride(0), _ptr(cu_seqlens), num_seqs, max_seqlen, q.shape[0], hq, hkv, d,
                                float(scale), _DT[q.dtype], _stream())
    _lib.check(rc, "llmss_attn_prefill")
    _count()


def attn_prefill_paged(q_view, k_cache, v_cache, block_tables, seq_lens, out_view, cu_q, num_seqs, max_q, hq, hkv, d, scale):
    """Multi-token queries over prior context: K / V (this step's included) are read from the paged cache."""
    lib = _lib.load()
    rc = lib.llmss_attn_prefill_paged(_ptr(q_view), q_view.stride(0), _ptr(k_cache), _ptr(v_cache), _ptr(block_tables),
                                      block_tables.shape[1], _ptr(seq_lens), _ptr(out_view), out_view.stride(0), _ptr(cu_q),
                                      num_seqs, max_q, q_view.shape[0], hq, hkv, d, k_cache.shape[2], k_cache.shape[0],
                                      float(scale), _DT[q_view.dtype], _stream())
    _lib.check(rc, "llmss_attn_prefill_paged")
    _count()


# --------------------------------------------------------------------------- embed / sample
def embed(tokens, wte, positions=None, wpe=None, pos_offset: int = 0):
    lib = _lib.load()
    T = tokens.numel()
    V, H = wte.shape
    out = torch.empty(T, H, dtype=wte.dtype, device=wte.device)
    rc = lib.llmss_embed(_ptr(tokens), _ptr(positions), _ptr(wte), _ptr(wpe), _ptr(out), T, H, V,
                         wpe.shape[0] if wpe is not None else 0, pos_offset, out.stride(0), _DT[wte.dtype], _stream())
    _lib.check(rc, "llmss_embed")
    _count()
    return out


def sample(logits, greedy, temperature, top_k, top_p, rng_state, vocab=None, out=None):
    lib = _lib.load()
    B, Vp = logits.shape
    assert logits.dtype == torch.float32 and logits.stride(1) == 1
    if out is None:
        out = torch.empty(B, dtype=torch.int32, device=logits.device)
    o32 = _ptr(out) if out.dtype == torch.int32 else None
    o64 = _ptr(out) if out.dtype == torch.int64 else None
    rc = lib.llmss_sample(_ptr(logits), logits.stride(0), B, vocab or Vp, _ptr(greedy), _ptr(temperature), _ptr(top_k),
                          _ptr(top_p), _ptr(rng_state), o32, o64, _stream())
    _lib.check(rc, "llmss_sample")
    _count()
    return out


def greedy_vocab_parallel(local_logits, symm, vocab: int, out=None):
    """K11 (greedy case): argmax over vocab-sharded fp32 logits [B, V/t] without gathering them.  Every rank
    reduces its shard to one 16-byte (value, global index) candidate per row, the candidates are gathered
    column-wise and merged identically on every rank."""
    lib = _lib.load()
    B, Vb = local_logits.shape
    assert local_logits.dtype == torch.float32 and local_logits.stride(1) == 1
    off = symm.rank * Vb
    n_valid = max(0, min(Vb, vocab - off))
    cand = torch.empty(B, 4, dtype=torch.int32, device=local_logits.device)
    _lib.check(lib.llmss_argmax_local(_ptr(local_logits), local_logits.stride(0), B, n_valid, off, _ptr(cand), _stream()),
               "llmss_argmax_local")
    allc = symm.all_gather_cols(cand)  # [B, t * 4]
    if out is None:
        out = torch.empty(B, dtype=torch.int32, device=local_logits.device)
    o32 = _ptr(out) if out.dtype == torch.int32 else None
    o64 = _ptr(out) if out.dtype == torch.int64 else None
    _lib.check(lib.llmss_argmax_merge(_ptr(allc), symm.nranks, B, o32, o64, _stream()), "llmss_argmax_merge")
    _count(2)
    return out


VP_KS = 128  # candidates a rank contributes per row: exact for 1 <= top_k <= 128 (8 ranks x 128 = the merge's 1024 slots)


def vp_summary(local_logits, n_valid: int, idx_offset: int, greedy, top_k, ks: int = VP_KS, temperature=None):
    """Per-row record of this rank's shard: [max, sum-exp, count, 0, vals[ks], idx[ks]] (fp32 / int32 bits)."""
    lib = _lib.load()
    B = local_logits.shape[0]
    assert local_logits.dtype == torch.float32 and local_logits.stride(1) == 1
    rec = torch.empty(B, 4 + 2 * ks, dtype=torch.float32, device=local_logits.device)
    _lib.check(lib.llmss_vp_summary(_ptr(local_logits), local_logits.stride(0), B, n_valid, idx_offset, _ptr(greedy),
                                    _ptr(top_k), _ptr(temperature), _ptr(rec), ks, _stream()), "llmss_vp_summary")
    _count()
    return rec


def vp_merge_sample(recs_all, nranks: int, vocab: int, greedy, temperature, top_k, top_p, rng_state, out=None,
                    ks: int = VP_KS):
    """recs_all: [B, nranks * (4 + 2 ks)] column-gathered records -> tokens [B]; advances rng_state."""
    lib = _lib.load()
    B = recs_all.shape[0]
    assert recs_all.is_contiguous() and recs_all.shape[1] == nranks * (4 + 2 * ks)
    if out is None:
        out = torch.empty(B, dtype=torch.int32, device=recs_all.device)
    o32 = _ptr(out) if out.dtype == torch.int32 else None
    o64 = _ptr(out) if out.dtype == torch.int64 else None
    _lib.check(lib.llmss_vp_merge_sample(_ptr(recs_all), nranks, ks, B, vocab, _ptr(greedy), _ptr(temperature),
                                         _ptr(top_k), _ptr(top_p), _ptr(rng_state), o32, o64, _stream()),
               "llmss_vp_merge_sample")
    _count()
    return out


def vp_multinomial_pick(local_logits, n_valid: int, idx_offset: int, recs_all, nranks: int, rank: int, greedy, temperature,
                        top_k, top_p, rng_state, ks: int = VP_KS):
    """Whole-distribution rows (top_k = 0, top_p = 1): the rank whose CDF interval holds the draw names the token; every
    other rank (and every other kind of row) answers -1.  Returns [B, 4] int32 (16 bytes per row for the gather)."""
    lib = _lib.load()
    B = local_logits.shape[0]
    cand = torch.empty(B, 4, dtype=torch.int32, device=local_logits.device)
    _lib.check(lib.llmss_vp_multinomial_pick(_ptr(local_logits), local_logits.stride(0), B, n_valid, idx_offset, _ptr(recs_all),
                                             nranks, rank, ks, _ptr(greedy), _ptr(temperature), _ptr(top_k), _ptr(top_p),
                                             _ptr(rng_state), _ptr(cand), _stream()), "llmss_vp_multinomial_pick")
    _count()
    return cand


def vp_multinomial_merge(cand_all, nranks: int, out):
    lib = _lib.load()
    o32 = _ptr(out) if out.dtype == torch.int32 else None
    o64 = _ptr(out) if out.dtype == torch.int64 else None
    _lib.check(lib.llmss_vp_multinomial_merge(_ptr(cand_all), nranks, out.shape[0], o32, o64, _stream()),
               "llmss_vp_multinomial_merge")
    _count()
    return out


def sample_vocab_parallel(local_logits, symm, vocab: int, greedy, temperature, top_k, top_p, rng_state, out=None,
                          whole_distribution_rows: bool = False):
    """K11: top-k / top-p / temperature sampling over vocab-sharded logits [B, V/t] without gathering them
    (reference path: list all_gather of [B, S, V/t] at layers.py:131-135 + rank-0 warpers + broadcast,
    generate.py:109-144).  Three launches: shard summary, 1 KiB-per-row record gather, identical merge + draw; batches
    with whole-distribution rows (top_k = 0, top_p = 1) add the owner pick, a 16-byte-per-row gather and its merge."""
    B, Vb = local_logits.shape
    off = symm.rank * Vb
    n_valid = max(0, min(Vb, vocab - off))
    rec = vp_summary(local_logits, n_valid, off, greedy, top_k, temperature=temperature)
    allr = symm.all_gather_cols(rec)
    cand_all = None
    if whole_distribution_rows:
        cand = vp_multinomial_pick(local_logits, n_valid, off, allr, symm.nranks, symm.rank, greedy, temperature, top_k, top_p,
                                   rng_state)
        cand_all = symm.all_gather_cols(cand)
    out = vp_merge_sample(allr, symm.nranks, vocab, greedy, temperature, top_k, top_p, rng_state, out)
    if cand_all is not None:
        vp_multinomial_merge(cand_all, symm.nranks, out)
    return out
