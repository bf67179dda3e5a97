"""Op front door: CPU tensors run the fp32 PyTorch oracles (ops.reference), CUDA tensors run
the hand-written sm_100a kernels (ops.cuda_impl -> libllmss_b200.so).  There is no silent
fallback on CUDA: a missing library or an unsupported shape raises.

The only library (non-hand-written) kernel still reachable on the CUDA path is the
``generic`` attention fallback used for prefill-with-prior-context (chunked prefill /
HF-compat forward with a multi-token query and a non-empty cache); see attention().
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import reference as ref
from .reference import (ACT_GELU_ERF, ACT_GELU_TANH, ACT_IDS, ACT_NONE, ACT_RELU, ACT_SILU, NORM_LAYER, NORM_RMS,
                        ROPE_GPTJ, ROPE_NEOX, ROPE_NONE)


def _cuda(t: torch.Tensor) -> bool:
    return t.is_cuda


_WARNED = set()


def _warn_once(msg: str):
    if msg not in _WARNED:
        _WARNED.add(msg)
        from ..utils.logging import logger

        logger.warning(msg)


_FUSED_PUSH_MAX = int(os.environ.get("LLMSS_FUSED_AR_MAX_PUSH", str(2 << 20)))
_TWOSHOT_MIN = int(os.environ.get("LLMSS_TWOSHOT_MIN_BYTES", str(256 << 10)))
_FUSED_TWOSHOT_MAX = int(os.environ.get("LLMSS_FUSED_TWOSHOT_MAX_BYTES", str(8 << 20)))


def _fused_ar_mode(cu, symm, M: int, N: int, K: int, esz: int) -> int:
    """How a row-parallel GEMM on CUDA performs its all-reduce: 0 = plain GEMM followed by the standalone
    collective, 5 = one-shot Lamport push in the epilogue, 6 = two-shot Lamport exchange in the epilogue
    (reduce-scatter by peer stores, result broadcast by one NVSwitch multicast store).

    Only the cluster split-K kernel fuses (each CTA owns a slice of one tile); shapes that run on the stream-K
    kernel measured slower fused than unfused.  The one-shot push is one hop but puts (t-1) x the message on
    the wire, the two-shot form two hops and ~1 x: from 4 ranks and `LLMSS_TWOSHOT_MIN_BYTES` upwards the
    latter wins (profiles/bench_gemm_ar.md).  The flag-synchronised in-switch reduction (mode 4,
    multimem.ld_reduce) costs ~15 us of latency and is left to the standalone collective for prefill sizes."""
    plan = cu.gemm_plan(M, N, K, symm.nranks)
    if plan.split <= 0 or plan.tiles > 2048:
        return 0
    msg = M * N * esz
    padded = plan.tiles * 128 * plan.bn * esz
    two_fits = padded <= _FUSED_TWOSHOT_MAX and (symm.nranks + 1) * padded <= symm.lam_bytes
    if two_fits and ((symm.nranks >= 4 and msg >= _TWOSHOT_MIN) or msg >= (1 << 20)):
        return 6   # measured: from 256 KB at 4+ ranks, from 1 MB even at 2 ranks (profiles/bench_gemm_ar.md)
    if plan.ar_bytes <= symm.lam_bytes and msg * (symm.nranks - 1) <= _FUSED_PUSH_MAX:
        return 5
    return 0


def linear(x, w, bias=None, act: int = ACT_NONE, gated: bool = False, act_start: int = 0, residual=None,
           out_dtype=None, comm=None):
    """y = epilogue(x @ w.T); with ``comm`` the product is summed over the TP group first
    (row-parallel layer: reference `TensorParallelRowLinear.forward`, layers.py:175-179) and
    bias/residual are applied once after the reduction."""
    reduce = comm is not None and not comm.single
    if _cuda(x):
        from . import cuda_impl as cu

        if not reduce:
            return cu.linear(x, w, bias, act, gated, act_start, residual, out_dtype)
        symm = comm.symm
        if symm is not None:
            mode = _fused_ar_mode(cu, symm, x.shape[0], w.shape[0], x.shape[1], x.element_size())
            if mode:
                return cu.linear(x, w, bias, ACT_NONE, False, 0, residual, None, symm=symm, collective=mode)
        y = cu.linear(x, w, None, ACT_NONE, False, 0, None, None)
        comm.all_reduce(y)
        if bias is not None:
            y = y + bias
        if residual is not None:
            y = y + residual
        return y
    if not reduce:
        return ref.linear(x, w, bias, act, gated, act_start, residual, out_dtype)
    y = ref.linear(x, w, None, ACT_NONE, False, 0, None, torch.float32)
    comm.all_reduce(y)
    if bias is not None:
        y = y + bias.float()
    if residual is not None:
        y = y + residual.float()
    return y.to(out_dtype or x.dtype)


def norm_residual(x, residual, weight, bias, eps: float, kind: int, x2=None):
    """Returns (norm(x [+x2] [+residual]), the pre-norm sum)."""
    if _cuda(x):
        from . import cuda_impl as cu

        return cu.norm_residual(x, residual, weight, bias, eps, kind, x2=x2)
    if x2 is not None:
        x = (x.float() + x2.float()).to(x.dtype)
    return ref.norm_residual(x, residual, weight, bias, eps, kind)


def rope_kv_write(qkv, positions, slot_mapping, k_cache, v_cache, hq, hkv, d, rotary_dim, style, base,
                  q_off, k_off, v_off):
    if _cuda(qkv):
        from . import cuda_impl as cu

        return cu.rope_kv_write(qkv, positions, slot_mapping, k_cache, v_cache, hq, hkv, d, rotary_dim, style, base,
                                q_off, k_off, v_off)
    return ref.rope_kv_write(qkv, positions, slot_mapping, k_cache, v_cache, hq, hkv, d, rotary_dim, style, base,
                             q_off, k_off, v_off)


def attention(qkv, out, batch, k_cache, v_cache, hq, hkv, d, scale, q_off, k_off, v_off, out_off=0):
    """Causal attention for a packed batch; writes ``out[:, out_off : out_off + hq*d]``.

    decode batches (one query per sequence) -> split-KV paged kernel (K9);
    fresh prefill (no prior context)        -> varlen causal prefill kernel (K8);
    multi-token queries over prior context  -> the same kernel in paged mode (K / V from the cache);
    unsupported head sizes / page sizes     -> generic gather path (oracle math).
    """
    T = qkv.shape[0]
    q_view = qkv[:, q_off:q_off + hq * d]
    out_view = out[:, out_off:out_off + hq * d]
    if _cuda(qkv):
        from . import cuda_impl as cu

        if d not in (64, 128, 256):
            _warn_once(f"head_dim={d} has no sm_100a attention kernel (64/128/256 do); using the generic PyTorch path")
        elif batch.is_decode:
            splits = batch.decode_splits or cu.decode_splits(batch.num_seqs, hq, hkv, d, batch.max_kv, qkv.dtype,
                                                                     k_cache.shape[2])
            cu.attn_decode(q_view, k_cache, v_cache, batch.block_tables, batch.seq_lens, out_view, hq, hkv, d, scale,
                           splits)
            return out
        elif batch.is_fresh_prefill and cu._lib.has("llmss_attn_prefill"):
            cu.attn_prefill(q_view, qkv[:, k_off:k_off + hkv * d], qkv[:, v_off:v_off + hkv * d], out_view,
                            batch.cu_q, batch.num_seqs, batch.max_q, hq, hkv, d, scale)
            return out
        elif (not batch.is_fresh_prefill and k_cache.shape[2] in (16, 32, 64) and k_cache.is_contiguous()
              and v_cache.is_contiguous() and cu._lib.has("llmss_attn_prefill_paged")):
            # multi-token queries over prior context (chunked prefill, HF-style forward with a past): same tcgen05
            # flash kernel, K / V tiles gathered page by page from the cache through the block table
            cu.attn_prefill_paged(q_view, k_cache, v_cache, batch.block_tables, batch.seq_lens, out_view, batch.cu_q,
                                  batch.num_seqs, batch.max_q, hq, hkv, d, scale)
            return out
        elif batch.is_fresh_prefill:
            # LIBRARY FALLBACK (torch SDPA), only reached when the own prefill kernel is absent from
            # the loaded library or the head_dim is unsupported by it.
            _sdpa_prefill(qkv, out_view, batch, hq, hkv, d, scale, q_off, k_off, v_off)
            return out
    o = ref.attention_paged(q_view.reshape(T, hq, d), k_cache, v_cache, batch.block_tables, batch.seq_lens,
                            batch.cu_q, scale)
    out_view.copy_(o.reshape(T, hq * d))
    return out


def _sdpa_prefill(qkv, out_view, batch, hq, hkv, d, scale, q_off, k_off, v_off):
    import torch.nn.functional as F

    cu = batch.cu_q.tolist()
    for i in range(batch.num_seqs):
        a, b = cu[i], cu[i + 1]
        if b == a:
            continue
        q = qkv[a:b, q_off:q_off + hq * d].reshape(1, b - a, hq, d).transpose(1, 2)
        k = qkv[a:b, k_off:k_off + hkv * d].reshape(1, b - a, hkv, d).transpose(1, 2)
        v = qkv[a:b, v_off:v_off + hkv * d].reshape(1, b - a, hkv, d).transpose(1, 2)
        o = F.scaled_dot_product_attention(q, k, v, is_causal=True, scale=scale, enable_gqa=(hq != hkv))
        out_view[a:b] = o.transpose(1, 2).reshape(b - a, hq * d)


def embed(tokens, wte, positions=None, wpe=None, pos_offset: int = 0):
    if _cuda(wte):
        from . import cuda_impl as cu

        return cu.embed(tokens, wte, positions, wpe, pos_offset)
    return ref.embed(tokens, wte, positions, wpe, pos_offset)


def sample(logits, greedy, temperature, top_k, top_p, rng_state, vocab: Optional[int] = None, out=None):
    if _cuda(logits):
        from . import cuda_impl as cu

        return cu.sample(logits, greedy, temperature, top_k, top_p, rng_state, vocab, out)
    tok = ref.sample(logits, greedy, temperature, top_k, top_p, rng_state, vocab).to(torch.int32)
    if out is not None:
        out.copy_(tok)
        return out
    return tok
