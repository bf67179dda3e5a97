"""One tensor-parallel decoder graph for every supported family.

The reference ships two 600-900 line HF-derived nn.Module trees (gptj_modeling.py,
gpt_bigcode_modeling.py).  Here a family is an :class:`ArchSpec` plus a weight loader that
lays the checkpoint out *per rank, fused for the kernels*:

    wqkv  [ (hq/t + 2*hkv_r) * d , H ]   rows = q heads of this rank, then k, then v
    wo    [ H , hq/t * d ]               row-parallel  (all-reduce fused into the GEMM)
    wup   [ F/t (x2 interleaved if gated) , H ]
    wdown [ H , F/t ]                    row-parallel  (all-reduce fused into the GEMM)

For parallel-residual models (GPT-J: x + attn(ln x) + mlp(ln x), gptj_modeling.py:296-310)
the two branches share their input and their outputs are summed, so they are fused further
into *two* GEMMs and *one* all-reduce per layer (the reference does 5 GEMMs + 2 all-reduces):

    w1 = [wk; wv; wq; wup]  ->  y = [k | v | q | act(up)]       (activation on the tail only)
    attention writes its context over q                      ->  z = y[:, kv_dim:] = [ctx | act(up)]
    w2 = [wo | wdown]       ->  out = all_reduce(z @ w2^T) + b_down

Sharding rules are the reference's (weights.py:97-115, layers.py:138-214)
so any unsharded HF safetensors checkpoint loads unchanged.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import torch
from torch import nn

from .. import ops
from ..engine.batch import BatchInput
from ..ops import (ACT_IDS, ACT_NONE, NORM_LAYER, NORM_RMS, ROPE_GPTJ, ROPE_NEOX, ROPE_NONE)
from ..parallel.comm import comm_for
from ..utils.tracing import trace


@dataclass
class ArchSpec:
    model_type: str
    hidden: int
    n_layers: int
    n_heads: int
    n_kv_heads: int
    head_dim: int
    ffn: int
    vocab: int
    max_pos: int
    norm: int = NORM_LAYER
    eps: float = 1e-5
    rope: int = ROPE_NONE
    rotary_dim: int = 0
    rope_base: float = 10000.0
    learned_pos: bool = False
    pos_offset: int = 0
    act: int = ACT_NONE
    gated: bool = False
    parallel_residual: bool = False
    head_bias: bool = False
    tie_head: bool = False
    attn_scale: Optional[float] = None

    @property
    def scale(self) -> float:
        return self.attn_scale if self.attn_scale is not None else self.head_dim ** -0.5


class LayerWeights:
    """Plain tensor container (no nn.Module overhead on the hot path)."""

    __slots__ = ("ln1_w", "ln1_b", "ln2_w", "ln2_b", "wqkv", "bqkv", "wo", "bo", "wup", "bup", "wdown", "bdown",
                 "w1", "b1", "w2", "b2")

    def __init__(self):
        for s in self.__slots__:
            setattr(self, s, None)

    def tensors(self):
        return {s: getattr(self, s) for s in self.__slots__ if getattr(self, s) is not None}


class DecoderLM(nn.Module):
    """Family-agnostic tensor-parallel decoder operating on packed token batches."""

    def __init__(self, spec: ArchSpec, process_group, device, dtype, fuse_parallel: bool = True):
        super().__init__()
        self.spec = spec
        self.process_group = process_group
        self.comm = comm_for(process_group, device)
        self.tp = self.comm.size
        self.rank = self.comm.rank
        self.device = torch.device(device)
        self.dtype = dtype
        # Heads that do not divide the shard count (GPT-2 XL: 25 heads at TP=2; the reference raises,
        # gptj_modeling.py:78-83) are padded with null heads at the end: their q/k/v rows and out-proj columns are
        # zero, so they attend uniformly over zero values and add nothing; no masking is needed anywhere.
        self.hq = -(-spec.n_heads // self.tp)
        self.heads_padded = self.hq * self.tp
        if self.heads_padded != spec.n_heads and spec.n_kv_heads not in (spec.n_heads, 1):
            raise ValueError(f"`num_attention_heads` must be divisible by `num_shards` for grouped-query models (got "
                             f"`num_attention_heads`: {spec.n_heads}, `num_key_value_heads`: {spec.n_kv_heads} and "
                             f"`num_shards`: {self.tp})")
        # KV heads: shard when there are enough, else replicate (MQA rule, gpt_bigcode_modeling.py:148-155)
        if spec.n_kv_heads == spec.n_heads:
            self.hkv = self.hq                      # multi-head: K/V heads follow (and are padded like) the q heads
            self.kv_head_start = self.rank * self.hkv
        elif spec.n_kv_heads >= self.tp:
            if spec.n_kv_heads % self.tp != 0:
                raise ValueError(f"num_key_value_heads={spec.n_kv_heads} not divisible by {self.tp} shards")
            self.hkv = spec.n_kv_heads // self.tp
            self.kv_head_start = self.rank * self.hkv
        else:
            if self.tp % spec.n_kv_heads != 0:
                raise ValueError(f"{self.tp} shards not a multiple of num_key_value_heads={spec.n_kv_heads}")
            self.hkv = 1
            self.kv_head_start = self.rank // (self.tp // spec.n_kv_heads)
        if spec.ffn % self.tp != 0:
            raise ValueError(f"ffn size {spec.ffn} not divisible by {self.tp} shards")
        self.ffn_r = spec.ffn // self.tp
        self.d = spec.head_dim
        self.q_dim, self.kv_dim = self.hq * self.d, self.hkv * self.d
        self.fuse_parallel = bool(fuse_parallel and spec.parallel_residual)
        # sequence-parallel prefill dataflow (GEMM->reduce-scatter, all-gather->GEMM).  Opt-in for now
        # (LLMSS_SP_PREFILL=1): measured at TP=2 on GPT-J-6B it is still slower than the fused all-reduce
        # path (814 ms vs 608 ms for 64 x 1024-token prompts); see profiles/bench_results.md.
        import os
        self.sp_prefill = os.environ.get("LLMSS_SP_PREFILL", "0") == "1"
        self.vocab_blk = -(-spec.vocab // self.tp)
        self.layers: List[LayerWeights] = []
        self.wte = self.wpe = self.lnf_w = self.lnf_b = self.w_head = self.b_head = None

    # ------------------------------------------------------------------ weights
    def finalize(self):
        """Build the fused GPT-J style operands once all per-layer tensors are loaded."""
        if self.fuse_parallel:
            for L in self.layers:
                q, k, v = L.wqkv[:self.q_dim], L.wqkv[self.q_dim:self.q_dim + self.kv_dim], L.wqkv[self.q_dim + self.kv_dim:]
                L.w1 = torch.cat([k, v, q, L.wup], dim=0).contiguous()
                if L.bqkv is not None or L.bup is not None:
                    zq = torch.zeros(self.q_dim + 2 * self.kv_dim, dtype=self.dtype, device=self.device)
                    bq = L.bqkv if L.bqkv is not None else zq
                    bq = torch.cat([bq[self.q_dim:self.q_dim + self.kv_dim], bq[self.q_dim + self.kv_dim:], bq[:self.q_dim]])
                    bu = L.bup if L.bup is not None else torch.zeros(L.wup.shape[0], dtype=self.dtype, device=self.device)
                    L.b1 = torch.cat([bq, bu]).contiguous()
                L.w2 = torch.cat([L.wo, L.wdown], dim=1).contiguous()
                if L.bo is not None and L.bdown is not None:
                    L.b2 = (L.bo.float() + L.bdown.float()).to(self.dtype)
                else:
                    L.b2 = L.bo if L.bo is not None else L.bdown
                L.wqkv = L.bqkv = L.wo = L.bo = L.wup = L.bup = L.wdown = L.bdown = None
        return self

    def num_parameters_local(self) -> int:
        n = sum(t.numel() for L in self.layers for t in L.tensors().values())
        for t in (self.wte, self.wpe, self.lnf_w, self.lnf_b, self.w_head, self.b_head):
            if t is not None:
                n += t.numel()
        return n

    # ------------------------------------------------------------------ forward
    def _layer_sequential(self, L: LayerWeights, x, resid, x2, batch: BatchInput, kc, vc):
        s = self.spec
        hn, resid = ops.norm_residual(x, resid, L.ln1_w, L.ln1_b, s.eps, s.norm, x2=x2)
        qkv = ops.linear(hn, L.wqkv, L.bqkv)
        q_off, k_off, v_off = 0, self.q_dim, self.q_dim + self.kv_dim
        ops.rope_kv_write(qkv, batch.positions, batch.slot_mapping, kc, vc, self.hq, self.hkv, self.d, s.rotary_dim,
                          s.rope, s.rope_base, q_off, k_off, v_off)
        # prefill: zeros, not empty -- padding rows (inert tokens that round the batch up for TP) are never written by the
        # prefill attention kernel, and a NaN left there would reach real rows through the next layer's 0 x V products
        # (decode rows only ever attend to their own sequence's cache, so the graph keeps the cheaper empty)
        ctx = (torch.empty if batch.is_decode else torch.zeros)(qkv.shape[0], self.q_dim, dtype=qkv.dtype, device=qkv.device)
        ops.attention(qkv, ctx, batch, kc, vc, self.hq, self.hkv, self.d, s.scale, q_off, k_off, v_off)
        if s.parallel_residual:
            o = ops.linear(ctx, L.wo, L.bo, comm=self.comm)
            u = ops.linear(hn, L.wup, L.bup, act=s.act, gated=s.gated)
            dn = ops.linear(u, L.wdown, L.bdown, comm=self.comm)
            return o, resid, dn
        o = ops.linear(ctx, L.wo, L.bo, comm=self.comm)
        hn2, resid = ops.norm_residual(o, resid, L.ln2_w, L.ln2_b, s.eps, s.norm)
        u = ops.linear(hn2, L.wup, L.bup, act=s.act, gated=s.gated)
        dn = ops.linear(u, L.wdown, L.bdown, comm=self.comm)
        return dn, resid, None

    def _layer_fused_parallel(self, L: LayerWeights, x, resid, batch: BatchInput, kc, vc):
        s = self.spec
        with trace("norm_residual"):
            hn, resid = ops.norm_residual(x, resid, L.ln1_w, L.ln1_b, s.eps, s.norm)
        qkv_dim = self.q_dim + 2 * self.kv_dim
        with trace("gemm_qkv_up"):
            y = ops.linear(hn, L.w1, L.b1, act=s.act, act_start=qkv_dim)    # [k | v | q | act(up)]
        k_off, v_off, q_off = 0, self.kv_dim, 2 * self.kv_dim
        with trace("rope_kv_write"):
            ops.rope_kv_write(y, batch.positions, batch.slot_mapping, kc, vc, self.hq, self.hkv, self.d, s.rotary_dim,
                              s.rope, s.rope_base, q_off, k_off, v_off)
        with trace("attention"):
            ops.attention(y, y, batch, kc, vc, self.hq, self.hkv, self.d, s.scale, q_off, k_off, v_off, out_off=q_off)
        with trace("gemm_out_down_allreduce"):
            out = ops.linear(y[:, q_off:], L.w2, L.b2, comm=self.comm)       # one GEMM + one all-reduce
        return out, resid

    # ------------------------------------------------------------------ sequence-parallel prefill
    def _sp_ok(self, batch: BatchInput) -> bool:
        """Sequence-parallel dataflow (north-star items 1+2): GEMM -> reduce-scatter after the row-parallel
        layers, all-gather -> GEMM before the column-parallel ones, norms / residuals on T/t rows per rank.
        Used for fresh prefill on CUDA when the shards fit the symmetric buffers; decode keeps the
        latency-optimal fused all-reduce."""
        symm = self.comm.symm
        T = batch.tokens.numel()
        if symm is None or self.tp == 1 or batch.is_decode or not batch.is_fresh_prefill or not self.sp_prefill:
            return False
        if T % (self.tp * 16) != 0 or T < self.tp * 64:
            return False
        s = self.spec
        k_max = s.hidden  # every all-gathered operand is a [T, hidden] activation
        if T * k_max * 2 > symm.buf_bytes:  # gathered activation lives in the bulk buffer
            return False
        return T * s.hidden * 2 <= symm.lam_bytes  # reduce-scatter exchange: one padded [T, H] tile set per parity

    def _forward_hidden_sp(self, batch: BatchInput, kv) -> torch.Tensor:
        from ..ops import cuda_impl as cu

        s, symm = self.spec, self.comm.symm
        T = batch.tokens.numel()
        rows = T // self.tp
        lo, hi = self.rank * rows, (self.rank + 1) * rows
        x = ops.embed(batch.tokens[lo:hi], self.wte, batch.positions[lo:hi] if s.learned_pos else None,
                      self.wpe if s.learned_pos else None, s.pos_offset)                   # [T/t, H] shard
        resid = None
        for li, L in enumerate(self.layers):
            kc, vc = kv.k(li), kv.v(li)
            if self.fuse_parallel:
                with trace("norm_residual"):
                    hn, resid = ops.norm_residual(x, resid, L.ln1_w, L.ln1_b, s.eps, s.norm)
                qkv_dim = self.q_dim + 2 * self.kv_dim
                with trace("allgather_gemm_qkv_up"):
                    y = cu.allgather_linear(hn, L.w1, symm, bias=L.b1, act=s.act, act_start=qkv_dim)
                k_off, v_off, q_off = 0, self.kv_dim, 2 * self.kv_dim
                ops.rope_kv_write(y, batch.positions, batch.slot_mapping, kc, vc, self.hq, self.hkv, self.d,
                                  s.rotary_dim, s.rope, s.rope_base, q_off, k_off, v_off)
                with trace("attention"):
                    ops.attention(y, y, batch, kc, vc, self.hq, self.hkv, self.d, s.scale, q_off, k_off, v_off,
                                  out_off=q_off)
                with trace("gemm_out_down_reduce_scatter"):
                    x = cu.linear_reduce_scatter(y[:, q_off:], L.w2, symm, bias=L.b2)
            else:
                hn, resid = ops.norm_residual(x, resid, L.ln1_w, L.ln1_b, s.eps, s.norm)
                qkv = cu.allgather_linear(hn, L.wqkv, symm, bias=L.bqkv)
                q_off, k_off, v_off = 0, self.q_dim, self.q_dim + self.kv_dim
                ops.rope_kv_write(qkv, batch.positions, batch.slot_mapping, kc, vc, self.hq, self.hkv, self.d,
                                  s.rotary_dim, s.rope, s.rope_base, q_off, k_off, v_off)
                ctx = torch.zeros(T, self.q_dim, dtype=qkv.dtype, device=qkv.device)
                ops.attention(qkv, ctx, batch, kc, vc, self.hq, self.hkv, self.d, s.scale, q_off, k_off, v_off)
                if s.parallel_residual:
                    o = cu.linear_reduce_scatter(ctx, L.wo, symm, bias=L.bo)
                    u = cu.allgather_linear(hn, L.wup, symm, bias=L.bup, act=s.act, gated=s.gated)
                    dn = cu.linear_reduce_scatter(u, L.wdown, symm, bias=L.bdown)
                    x = o + dn
                else:
                    o = cu.linear_reduce_scatter(ctx, L.wo, symm, bias=L.bo)
                    hn2, resid = ops.norm_residual(o, resid, L.ln2_w, L.ln2_b, s.eps, s.norm)
                    u = cu.allgather_linear(hn2, L.wup, symm, bias=L.bup, act=s.act, gated=s.gated)
                    x = cu.linear_reduce_scatter(u, L.wdown, symm, bias=L.bdown)
        hn, _ = ops.norm_residual(x, resid, self.lnf_w, self.lnf_b, s.eps, s.norm)        # [T/t, H] shard
        return hn

    def _gather_last_rows(self, hn_shard: torch.Tensor, batch: BatchInput) -> torch.Tensor:
        """[B, H] last-token hidden states from the row shards: each rank fills the rows it owns, one
        small all-reduce makes them identical everywhere."""
        rows = hn_shard.shape[0]
        lo = self.rank * rows
        idx = batch.last_idx
        mine = (idx >= lo) & (idx < lo + rows)
        out = torch.zeros(idx.numel(), hn_shard.shape[1], dtype=hn_shard.dtype, device=hn_shard.device)
        sel = torch.nonzero(mine, as_tuple=False).flatten()
        if sel.numel():
            out[sel] = hn_shard.index_select(0, idx[sel] - lo)
        pad = (-out.numel() * out.element_size()) % 16
        self.comm.all_reduce(out if pad == 0 else out)  # [B, H] rows are 16-byte multiples (H % 8 == 0)
        return out

    def forward_hidden(self, batch: BatchInput, kv) -> torch.Tensor:
        """Final-normed hidden states for every packed token: [T, H]."""
        s = self.spec
        x = ops.embed(batch.tokens, self.wte, batch.positions if s.learned_pos else None,
                      self.wpe if s.learned_pos else None, s.pos_offset)
        resid, x2 = None, None
        for li, L in enumerate(self.layers):
            kc, vc = kv.k(li), kv.v(li)
            if self.fuse_parallel:
                x, resid = self._layer_fused_parallel(L, x, resid, batch, kc, vc)
            else:
                x, resid, x2 = self._layer_sequential(L, x, resid, x2, batch, kc, vc)
        hn, _ = ops.norm_residual(x, resid, self.lnf_w, self.lnf_b, s.eps, s.norm, x2=x2)
        return hn

    def logits(self, hidden_rows: torch.Tensor) -> torch.Tensor:
        """fp32 logits [R, tp * vocab_blk] (columns >= spec.vocab are padding).

        Vocab-parallel head GEMM on the selected rows only (the reference runs it on all S
        positions, quirk Q11) followed by one all-gather (reference N3, layers.py:131-135)."""
        with trace("lm_head"):
            local = ops.linear(hidden_rows, self.w_head, self.b_head, out_dtype=torch.float32)
        if self.tp == 1:
            return local
        with trace("logits_all_gather"):
            return self.comm.all_gather_cat(local, dim=-1)

    def local_logits(self, hidden_rows: torch.Tensor) -> torch.Tensor:
        """This rank's vocab shard of the fp32 logits, [R, vocab_blk] (no collective); the vocab-parallel
        greedy sampler (`ops.cuda_impl.greedy_vocab_parallel`) consumes it directly."""
        with trace("lm_head"):
            return ops.linear(hidden_rows, self.w_head, self.b_head, out_dtype=torch.float32)

    def forward(self, batch: BatchInput, kv, last_only: bool = True, gather_logits: bool = True) -> torch.Tensor:
        if last_only and self._sp_ok(batch):
            h = self._gather_last_rows(self._forward_hidden_sp(batch, kv), batch)
        else:
            h = self.forward_hidden(batch, kv)
            if last_only:
                h = h.index_select(0, batch.last_idx)
        return self.logits(h) if gather_logits else self.local_logits(h)


# ----------------------------------------------------------------------------- loading helpers
class ShardReader:
    """Reads the rank-local part of HF checkpoint tensors (lazy safetensors slices)."""

    def __init__(self, weights, model: DecoderLM):
        self.w = weights
        self.m = model
        self.rank, self.tp = model.rank, model.tp

    def full(self, name):
        return self.w.get_tensor(name)

    def maybe(self, name):
        return self.w.get_tensor(name) if self.w.has(name) else None

    def rows(self, name, start, stop, transposed=False):
        """Rows [start, stop) of the logical [out, in] weight (bias: elements)."""
        shape = self.w.get_shape(name)
        if len(shape) == 1:
            return self.w.get_range(name, 0, start, stop)
        if transposed:  # GPT-2 Conv1D stores [in, out]
            return self.w.get_range(name, 1, start, stop).t().contiguous()
        return self.w.get_range(name, 0, start, stop)

    def cols(self, name, start, stop, transposed=False):
        """Input-dim range [start, stop) of the logical [out, in] weight."""
        if transposed:
            return self.w.get_range(name, 0, start, stop).t().contiguous()
        return self.w.get_range(name, 1, start, stop).contiguous()

    def _pad(self, t, n, dim):
        if t.shape[dim] == n:
            return t
        shape = list(t.shape)
        shape[dim] = n - t.shape[dim]
        return torch.cat([t, torch.zeros(shape, dtype=t.dtype, device=t.device)], dim=dim).contiguous()

    def rows_padded(self, name, start, stop, limit, transposed=False):
        """rows() for head-sharded projections: rows at or beyond ``limit`` (the end of the real heads) are zero."""
        shape = self.w.get_shape(name)
        other = None if len(shape) == 1 else (shape[0] if transposed else shape[1])
        hi = min(stop, limit)
        if start >= hi:
            z = (stop - start,) if other is None else (stop - start, other)
            return torch.zeros(z, dtype=self.m.dtype, device=self.m.device)
        return self._pad(self.rows(name, start, hi, transposed), stop - start, 0)

    def cols_padded(self, name, start, stop, limit, transposed=False):
        """cols() for the attention output projection: input columns of null heads are zero."""
        shape = self.w.get_shape(name)
        out = shape[1] if transposed else shape[0]
        hi = min(stop, limit)
        if start >= hi:
            return torch.zeros(out, stop - start, dtype=self.m.dtype, device=self.m.device)
        return self._pad(self.cols(name, start, hi, transposed), stop - start, 1)

    def q_rows(self, name, base=0, transposed=False):
        """This rank's q-head rows of a projection whose q block starts at row ``base``."""
        m = self.m
        lo = base + m.rank * m.q_dim
        return self.rows_padded(name, lo, lo + m.q_dim, base + m.spec.n_heads * m.d, transposed)

    def kv_rows(self, name, base=0, transposed=False):
        """This rank's k (or v) head rows of a projection whose k (v) block starts at row ``base``."""
        m = self.m
        lo = base + m.kv_head_start * m.d
        return self.rows_padded(name, lo, lo + m.kv_dim, base + m.spec.n_kv_heads * m.d, transposed)

    def o_cols(self, name, transposed=False):
        m = self.m
        lo = m.rank * m.q_dim
        return self.cols_padded(name, lo, lo + m.q_dim, m.spec.n_heads * m.d, transposed)

    def shard_rows(self, name, total, transposed=False):
        blk = total // self.tp
        return self.rows(name, self.rank * blk, (self.rank + 1) * blk, transposed)

    def shard_cols(self, name, total, transposed=False):
        blk = total // self.tp
        return self.cols(name, self.rank * blk, (self.rank + 1) * blk, transposed)

    def head(self, name, bias_name=None):
        """Vocab-parallel LM head block, ceil-divided and zero padded (fixes quirk Q10)."""
        m = self.m
        V, blk = m.spec.vocab, m.vocab_blk
        lo, hi = min(V, self.rank * blk), min(V, (self.rank + 1) * blk)
        w = self.w.get_range(name, 0, lo, hi)
        if hi - lo < blk:
            w = torch.cat([w, torch.zeros(blk - (hi - lo), w.shape[1], dtype=w.dtype, device=w.device)])
        b = None
        if bias_name is not None and self.w.has(bias_name):
            b = self.w.get_range(bias_name, 0, lo, hi)
            if hi - lo < blk:
                b = torch.cat([b, torch.zeros(blk - (hi - lo), dtype=b.dtype, device=b.device)])
        return w.contiguous(), b


def interleave_gate_up(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """rows (g0, u0, g1, u1, ...) so that the GEMM epilogue can form act(g) * u lane-locally."""
    return torch.stack([gate, up], dim=1).reshape(gate.shape[0] * 2, *gate.shape[1:]).contiguous()
