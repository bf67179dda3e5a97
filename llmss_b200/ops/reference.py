"""Plain-PyTorch fp32 oracles for every op in ``llmss_b200.ops``.

These are (a) the CPU execution path (BASELINE config 1: generate.py on CPU, TP=1; gloo
multi-process tests) and (b) the numerical reference every CUDA kernel in ``csrc/`` is
tested against (tests/test_gpu_*.py).  They mirror the math of the reference's ATen call
sites listed in SURVEY.md section 2.6 (G1-G16), not its code.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

ACT_NONE, ACT_GELU_TANH, ACT_GELU_ERF, ACT_RELU, ACT_SILU = 0, 1, 2, 3, 4
ACT_IDS = {
    None: ACT_NONE, "none": ACT_NONE,
    "gelu_new": ACT_GELU_TANH, "gelu_pytorch_tanh": ACT_GELU_TANH, "gelu_fast": ACT_GELU_TANH,
    "gelu": ACT_GELU_ERF, "relu": ACT_RELU, "silu": ACT_SILU, "swish": ACT_SILU,
}
NORM_LAYER, NORM_RMS = 0, 1
ROPE_NONE, ROPE_GPTJ, ROPE_NEOX = 0, 1, 2


def apply_act(x: torch.Tensor, act: int) -> torch.Tensor:
    if act == ACT_NONE:
        return x
    if act == ACT_GELU_TANH:  # gelu_new == gelu_pytorch_tanh (gptj_modeling.py:266, gpt_bigcode_modeling.py:321)
        return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * x * x * x)))
    if act == ACT_GELU_ERF:
        return F.gelu(x)
    if act == ACT_RELU:
        return F.relu(x)
    if act == ACT_SILU:
        return F.silu(x)
    raise ValueError(act)


def norm_residual(x, residual, weight, bias, eps: float, kind: int, out_dtype=None):
    """(x [+ residual]) -> (norm(sum), sum).  G2 + G13 fused (and the G18 capability)."""
    s = x.float() if residual is None else x.float() + residual.float()
    if kind == NORM_LAYER:
        y = F.layer_norm(s, (s.shape[-1],), weight.float(), None if bias is None else bias.float(), eps)
    else:
        var = s.pow(2).mean(-1, keepdim=True)
        y = s * torch.rsqrt(var + eps) * weight.float()
    od = out_dtype or x.dtype
    return y.to(od), s.to(od)


def linear(x, w, bias=None, act: int = ACT_NONE, gated: bool = False, act_start: int = 0,
           residual=None, out_dtype=None):
    """y = act(x @ w.T + bias) (+ residual).  ``w`` is [N, K] (HF nn.Linear layout).

    ``gated``: rows of w are interleaved (gate_0, up_0, gate_1, up_1, ...) and the result is
    act(gate) * up with N/2 columns (Llama SwiGLU fused into the epilogue).
    ``act_start``: activation applies to output features >= act_start only (GPT-J fused
    QKV + fc_in GEMM).
    """
    y = x.float() @ w.float().t()
    if bias is not None:
        y = y + bias.float()
    if gated:
        y = apply_act(y[..., 0::2], act) * y[..., 1::2]
    elif act != ACT_NONE:
        if act_start > 0:
            y = torch.cat([y[..., :act_start], apply_act(y[..., act_start:], act)], dim=-1)
        else:
            y = apply_act(y, act)
    if residual is not None:
        y = y + residual.float()
    return y.to(out_dtype or x.dtype)


def rope_cos_sin(positions: torch.Tensor, rotary_dim: int, base: float):
    inv_freq = 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float32, device=positions.device) / rotary_dim))
    ang = positions.float()[:, None] * inv_freq[None, :]
    return torch.cos(ang), torch.sin(ang)  # [T, rotary_dim/2]


def apply_rope(x: torch.Tensor, cos, sin, rotary_dim: int, style: int) -> torch.Tensor:
    """x: [T, heads, d] fp32.  GPT-J = interleaved pairs (gptj_modeling.py:37-47); NeoX/Llama = half split."""
    if style == ROPE_NONE or rotary_dim == 0:
        return x
    rot, rest = x[..., :rotary_dim], x[..., rotary_dim:]
    c, s = cos[:, None, :], sin[:, None, :]
    if style == ROPE_GPTJ:
        x1, x2 = rot[..., 0::2], rot[..., 1::2]
        o1, o2 = x1 * c - x2 * s, x2 * c + x1 * s
        rot = torch.stack((o1, o2), dim=-1).flatten(-2)
    else:
        h = rotary_dim // 2
        x1, x2 = rot[..., :h], rot[..., h:]
        rot = torch.cat((x1 * c - x2 * s, x2 * c + x1 * s), dim=-1)
    return torch.cat((rot, rest), dim=-1)


def rope_kv_write(qkv, positions, slot_mapping, k_cache, v_cache, hq: int, hkv: int, d: int,
                  rotary_dim: int, style: int, base: float, q_off: int, k_off: int, v_off: int):
    """RoPE on q,k in place inside ``qkv`` rows; write k,v to the paged cache (G5 + G6).

    qkv: [T, W] (row may be wider than the qkv columns); caches: [num_blocks, hkv, bs, d].
    slot = block * bs + offset; slot < 0 skips the write (padding rows of a captured graph).
    """
    T = qkv.shape[0]
    q = qkv[:, q_off:q_off + hq * d].reshape(T, hq, d).float()
    k = qkv[:, k_off:k_off + hkv * d].reshape(T, hkv, d).float()
    v = qkv[:, v_off:v_off + hkv * d].reshape(T, hkv, d)
    if style != ROPE_NONE and rotary_dim > 0:
        cos, sin = rope_cos_sin(positions, rotary_dim, base)
        q = apply_rope(q, cos, sin, rotary_dim, style)
        k = apply_rope(k, cos, sin, rotary_dim, style)
        qkv[:, q_off:q_off + hq * d] = q.reshape(T, hq * d).to(qkv.dtype)
        qkv[:, k_off:k_off + hkv * d] = k.reshape(T, hkv * d).to(qkv.dtype)
    bs = k_cache.shape[2]
    valid = slot_mapping >= 0
    slots = slot_mapping[valid].long()
    blk, off = slots // bs, slots % bs
    k_cache[blk, :, off, :] = k[valid].to(k_cache.dtype)
    v_cache[blk, :, off, :] = v[valid].to(v_cache.dtype)


def attention_paged(q, k_cache, v_cache, block_tables, seq_lens, cu_q, scale: float):
    """Causal attention of packed queries against the paged cache (G7/G8 without S^2 masks).

    q: [T, hq, d]; sequence i owns query rows cu_q[i]:cu_q[i+1], which are the *last*
    q_len_i positions of its seq_lens[i] cached tokens.  Returns [T, hq, d] in q.dtype.
    """
    T, hq, d = q.shape
    hkv, bs = k_cache.shape[1], k_cache.shape[2]
    g = hq // hkv
    out = torch.zeros(T, hq, d, dtype=torch.float32, device=q.device)
    cu = cu_q.tolist()
    lens = seq_lens.tolist()
    for i in range(len(lens)):
        q0, q1 = cu[i], cu[i + 1]
        L, ql = lens[i], q1 - q0
        if ql == 0 or L == 0:
            continue
        nb = (L + bs - 1) // bs
        blocks = block_tables[i, :nb].long()
        k = k_cache[blocks].permute(1, 0, 2, 3).reshape(hkv, nb * bs, d)[:, :L].float()
        v = v_cache[blocks].permute(1, 0, 2, 3).reshape(hkv, nb * bs, d)[:, :L].float()
        qi = q[q0:q1].float().permute(1, 0, 2)  # [hq, ql, d]
        k = k.repeat_interleave(g, dim=0)
        v = v.repeat_interleave(g, dim=0)
        s = torch.matmul(qi, k.transpose(1, 2)) * scale  # [hq, ql, L]
        qpos = torch.arange(L - ql, L, device=q.device)[:, None]
        kpos = torch.arange(L, device=q.device)[None, :]
        s = s.masked_fill(kpos > qpos, float("-inf"))
        p = torch.softmax(s, dim=-1)
        out[q0:q1] = torch.matmul(p, v).permute(1, 0, 2)
    return out.to(q.dtype)


def embed(tokens, wte, positions=None, wpe=None, pos_offset: int = 0, out_dtype=None):
    """Token (+ learned position) embedding (G1; replicated tables, no collective)."""
    h = F.embedding(tokens.long(), wte)
    if wpe is not None:
        h = h + F.embedding(positions.long() + pos_offset, wpe)
    return h.to(out_dtype or wte.dtype)


def _fmix32(x):
    x = x ^ (x >> 16)
    x = (x * 0x85EBCA6B) & 0xFFFFFFFF
    x = x ^ (x >> 13)
    x = (x * 0xC2B2AE35) & 0xFFFFFFFF
    return x ^ (x >> 16)


def rng_uniform(seed: torch.Tensor, step: torch.Tensor) -> torch.Tensor:
    """Counter-based uniforms in [0,1), one per row, keyed on the request's (seed, step) only (not on the
    batch row: a request samples the same tokens in any slot); bit-identical to csrc/sample.cu."""
    s, t = seed.long() & 0xFFFFFFFF, step.long() & 0xFFFFFFFF
    x = (0x9E3779B1 + t * 0x85EBCA77 + s * 0xC2B2AE3D) & 0xFFFFFFFF
    x = _fmix32(_fmix32(x))
    return (x >> 8).float() * (1.0 / 16777216.0)


MAX_CANDIDATES = 1024


def sample(logits, greedy, temperature, top_k, top_p, rng_state, vocab: Optional[int] = None,
           u_shift: float = 0.0) -> torch.Tensor:
    """Next-token choice per row (G15/G16 with the *intended* semantics).

    Order = the reference's warper list order (generate.py:114-120): top-p on the raw
    distribution, then top-k, then temperature; then one inverse-CDF draw over the
    survivors in (probability desc, index asc) order.  top_k == 0 means "off": with
    top_p < 1 the nucleus is searched among the 1024 most probable tokens; with
    top_p >= 1 it is a plain multinomial in index order.  ``rng_state`` is int32/int64
    [B, 2] = (seed, step) per row and its step column is advanced.  All ranks hold
    identical logits and rng state, so they agree without a broadcast (vs. N6).
    """
    B, V = logits.shape
    x = logits.float().clone()
    if vocab is not None and vocab < V:
        x[:, vocab:] = float("-inf")
        V = vocab
        x = x[:, :V]
    greedy = greedy.bool()
    arg = x.argmax(-1)
    # u_shift nudges the uniform draw (tests use it to find the rows whose choice sits on a CDF boundary)
    u = (rng_uniform(rng_state[:, 0], rng_state[:, 1]).double()[:, None] + u_shift).clamp(0.0, 1.0 - 2.0 ** -24)
    rng_state[:, 1] += 1
    xmax = x.max(-1, keepdim=True).values
    temp = temperature.float()[:, None]

    # candidate path
    sl, si = torch.sort(x, dim=-1, descending=True, stable=True)
    sp = torch.exp(sl - xmax) / torch.exp(x - xmax).sum(-1, keepdim=True)
    prev = torch.cumsum(sp, dim=-1) - sp
    k_eff = torch.where((top_k > 0) & (top_k <= MAX_CANDIDATES), top_k, torch.full_like(top_k, MAX_CANDIDATES))
    k_eff = k_eff.clamp(max=V).long()
    thr_k = torch.gather(sl, 1, (k_eff - 1)[:, None])
    cand = sl >= thr_k
    n_p = (cand & (prev < top_p.float()[:, None])).sum(-1).clamp(min=1)
    thr_p = torch.gather(sl, 1, (n_p - 1)[:, None])
    keep = cand & (sl >= thr_p)
    w = torch.where(keep, torch.exp((sl - xmax) / temp), torch.zeros_like(sl)).double()
    c = torch.cumsum(w, dim=-1)
    pick = (c <= u * c[:, -1:]).sum(-1)
    pick = torch.minimum(pick, keep.sum(-1) - 1)
    tok_cand = torch.gather(si, 1, pick[:, None])[:, 0]

    # plain multinomial path (index order)
    wf = torch.exp((x - xmax) / temp).double()
    cf = torch.cumsum(wf, dim=-1)
    tok_full = (cf <= u * cf[:, -1:]).sum(-1).clamp(max=V - 1)

    plain = (top_k <= 0) & (top_p.float() >= 1.0)
    tok = torch.where(plain, tok_full, tok_cand)
    return torch.where(greedy, arg, tok)


# --------------------------------------------------------------------------- vocab-parallel sampling oracle
def shard_summary(local_logits, offset: int, vocab: int, k: int):
    """What one TP rank would contribute to a vocab-parallel sampler (K11, general case) for its
    shard ``local_logits`` [B, Vb] whose column 0 is global token ``offset``: the local maximum, and
    every local element >= the local k-th largest (value, global index) -- ties included, which is
    what makes the merge exact (an element >= the global k-th value is >= its rank's local k-th value).
    Columns at or beyond ``vocab`` are padding."""
    x = local_logits.float().clone()
    B, Vb = x.shape
    n_valid = max(0, min(Vb, vocab - offset))
    x[:, n_valid:] = float("-inf")
    out = []
    for b in range(B):
        row = x[b, :n_valid]
        if n_valid == 0:
            out.append(dict(max=float("-inf"), vals=row.new_zeros(0), idx=torch.zeros(0, dtype=torch.long), row=row,
                            offset=offset))
            continue
        kk = min(k, n_valid)
        thr = torch.topk(row, kk).values[-1]
        sel = (row >= thr).nonzero()[:, 0]
        out.append(dict(max=row.max().item(), vals=row[sel], idx=sel + offset, row=row, offset=offset))
    return out


def sample_vocab_parallel(shards, greedy, temperature, top_k, top_p, rng_state, vocab: int, k_summary: int = 64,
                          u_shift: float = 0.0):
    """`sample()` computed from per-rank shard summaries only (the algorithm a vocab-parallel CUDA sampler
    follows; replaces the [B, V] logits all-gather of layers.py:131-135 + generate.py:109-129).

    ``shards``: list over ranks of fp32 [B, Vb] local logits (rank r starts at token r * Vb).
    Exchanged per row and rank: local max, local sum-exp, the local top-``k_summary`` candidates
    (+ ties); for the plain multinomial path two scalars (weight sum, count below the target).
    Rows whose effective k exceeds ``k_summary`` (top_k = 0 with top_p < 1 searches 1024 candidates)
    are outside this path -- the caller gathers the logits for them."""
    t = len(shards)
    B, Vb = shards[0].shape
    summ = [shard_summary(s, r * Vb, vocab, k_summary) for r, s in enumerate(shards)]
    u = (rng_uniform(rng_state[:, 0], rng_state[:, 1]).double() + u_shift).clamp(0.0, 1.0 - 2.0 ** -24)
    rng_state[:, 1] += 1
    toks = torch.zeros(B, dtype=torch.long)
    for b in range(B):
        rows = [summ[r][b] for r in range(t)]
        M = max(r_["max"] for r_ in rows)
        vals = torch.cat([r_["vals"] for r_ in rows])
        idx = torch.cat([r_["idx"] for r_ in rows])
        # global order: probability descending, index ascending
        o = torch.argsort(idx, stable=True)
        vals, idx = vals[o], idx[o]
        o = torch.argsort(vals, descending=True, stable=True)
        sl, si = vals[o], idx[o]
        if bool(greedy[b]):
            toks[b] = si[0]
            continue
        T = float(temperature[b])
        k, p = int(top_k[b]), float(top_p[b])
        if k <= 0 and p >= 1.0:
            # plain multinomial in index order: per-rank weight sums, then per-rank counts below the target
            w = [torch.exp((r_["row"] - M) / T).double() for r_ in rows]
            sums = [x.sum() for x in w]
            total = sum(sums)
            target = u[b] * total
            count, prefix = 0, 0.0
            for r in range(t):
                if len(w[r]):
                    count += int(((torch.cumsum(w[r], 0) + prefix) <= target).sum())
                prefix = prefix + sums[r]
            toks[b] = min(count, vocab - 1)
            continue
        k_eff = min(k if 0 < k <= MAX_CANDIDATES else MAX_CANDIDATES, vocab)
        if k_eff > k_summary:
            raise ValueError(f"row {b}: effective top-k {k_eff} exceeds the summary size {k_summary}")
        # global sum-exp from the ranks' (local max, local sum-exp) pairs
        Z = sum(float(torch.exp(r_["row"] - r_["max"]).sum()) * math.exp(r_["max"] - M) for r_ in rows if len(r_["row"]))
        thr_k = sl[k_eff - 1]
        head = sl >= thr_k            # complete: everything >= the global k-th value is in the union
        sl, si = sl[head], si[head]
        sp = torch.exp(sl - M) / Z
        prev = torch.cumsum(sp, 0) - sp
        n_p = max(int((prev < p).sum()), 1)
        keep = sl >= sl[n_p - 1]
        wt = torch.where(keep, torch.exp((sl - M) / T), torch.zeros_like(sl)).double()
        c = torch.cumsum(wt, 0)
        pick = min(int((c <= u[b] * c[-1]).sum()), int(keep.sum()) - 1)
        toks[b] = si[pick]
    return toks
