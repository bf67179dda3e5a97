"""Baseline tier R2: "the reference's library path at equal shapes" for models the reference cannot load.

The reference registers only `gptj` and `gpt_bigcode` (custom_modeling/__init__.py:4-7), so BASELINE.json's
Llama-2 7B / 13B and OPT-1.3B configs have no reference model code to run.  What the reference *would* run for
them is its own tensor-parallel layer library, so this file assembles exactly those classes -- imported unmodified
from baseline/_ref: `TensorParallelEmbedding`, `TensorParallelColumnLinear`, `TensorParallelRowLinear`,
`TensorParallelHead` (layers.py:79-214), loaded through the reference's `Weights` -- into a decoder that follows
the structure of the reference's GPT-J model file: separate q / k / v / (gate / up) GEMMs through `F.linear`, the
row-parallel projections followed by a blocking NCCL all-reduce, KV cache grown with `torch.cat`
(gptj_modeling.py:229-233), attention as fp32 `matmul` + causal `where` + `softmax` + `matmul`
(gptj_modeling.py:128-169), ATen norms and activations, logits for all positions + list all-gather.
None of this framework's models, kernels or engine is on this path.  Every number produced with it is labelled
"reference library path at equal shapes (R2)", never "reference model".
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn


def _rope_neox(q, k, pos, base, d):
    inv = 1.0 / (base ** (torch.arange(0, d, 2, device=q.device, dtype=torch.float32) / d))
    ang = pos.float()[:, None] * inv[None, :]                  # [S, d/2]
    cos, sin = torch.cat([ang.cos(), ang.cos()], -1), torch.cat([ang.sin(), ang.sin()], -1)
    cos, sin = cos[None, None].to(q.dtype), sin[None, None].to(q.dtype)

    def rot(x):
        x1, x2 = x[..., : d // 2], x[..., d // 2:]
        return torch.cat([-x2, x1], -1)

    return q * cos + rot(q) * sin, k * cos + rot(k) * sin


class _Attn(nn.Module):
    """q / k / v column-parallel, output row-parallel (+ all-reduce); mirrors GPTJAttention.forward."""

    def __init__(self, config, prefix, weights, names, bias, heads, kv_heads, d, rope_base):
        super().__init__()
        from llmss.server.models.utils.layers import TensorParallelColumnLinear, TensorParallelRowLinear

        t = weights.process_group.size()
        assert heads % t == 0 and kv_heads % t == 0, "R2: heads must divide the shard count (the reference's own rule)"
        self.h, self.hkv, self.d, self.rope_base = heads // t, kv_heads // t, d, rope_base
        qn, kn, vn, on = names
        self.q_proj = TensorParallelColumnLinear.load(config, f"{prefix}.{qn}", weights, bias)
        self.k_proj = TensorParallelColumnLinear.load(config, f"{prefix}.{kn}", weights, bias)
        self.v_proj = TensorParallelColumnLinear.load(config, f"{prefix}.{vn}", weights, bias)
        self.o_proj = TensorParallelRowLinear.load(config, f"{prefix}.{on}", weights, bias)

    def forward(self, x, pos, past):
        B, S, _ = x.shape
        q = self.q_proj(x).view(B, S, self.h, self.d).transpose(1, 2)
        k = self.k_proj(x).view(B, S, self.hkv, self.d).transpose(1, 2)
        v = self.v_proj(x).view(B, S, self.hkv, self.d).transpose(1, 2)
        if self.rope_base is not None:
            q, k = _rope_neox(q, k, pos, self.rope_base, self.d)
        if past is not None:
            k = torch.cat((past[0], k), dim=-2)
            v = torch.cat((past[1], v), dim=-2)
        present = (k, v)
        T = k.shape[-2]
        if self.hkv != self.h:
            k = k.repeat_interleave(self.h // self.hkv, dim=1)
            v = v.repeat_interleave(self.h // self.hkv, dim=1)
        # fp32 scores, causal where-mask, softmax, cast back: gptj_modeling.py:141-167
        w = torch.matmul(q.float(), k.float().transpose(-1, -2)) / (self.d ** 0.5)
        causal = torch.ones(T, T, dtype=torch.bool, device=x.device).tril()[T - S:T, :T]
        w = torch.where(causal, w, torch.tensor(torch.finfo(w.dtype).min, dtype=w.dtype, device=w.device))
        w = F.softmax(w, dim=-1).to(v.dtype)
        o = torch.matmul(w, v).transpose(1, 2).reshape(B, S, self.h * self.d)
        return self.o_proj(o), present


class _RMSNorm(nn.Module):
    def __init__(self, weight, eps):
        super().__init__()
        self.weight, self.eps = nn.Parameter(weight), eps

    def forward(self, x):
        v = x.float()
        v = v * torch.rsqrt(v.pow(2).mean(-1, keepdim=True) + self.eps)
        return (self.weight * v.to(x.dtype))


class R2Llama(nn.Module):
    def __init__(self, config, weights):
        super().__init__()
        from llmss.server.models.utils.layers import (TensorParallelColumnLinear, TensorParallelEmbedding,
                                                      TensorParallelHead, TensorParallelRowLinear)

        c = config
        d = c.hidden_size // c.num_attention_heads
        self.embed = TensorParallelEmbedding("model.embed_tokens", weights)
        self.layers = nn.ModuleList()
        for i in range(c.num_hidden_layers):
            p = f"model.layers.{i}"
            L = nn.Module()
            L.ln1 = _RMSNorm(weights.get_tensor(f"{p}.input_layernorm.weight"), c.rms_norm_eps)
            L.ln2 = _RMSNorm(weights.get_tensor(f"{p}.post_attention_layernorm.weight"), c.rms_norm_eps)
            L.attn = _Attn(c, f"{p}.self_attn", weights, ("q_proj", "k_proj", "v_proj", "o_proj"), False,
                           c.num_attention_heads, c.num_key_value_heads, d, float(getattr(c, "rope_theta", 10000.0) or 10000.0))
            L.gate = TensorParallelColumnLinear.load(c, f"{p}.mlp.gate_proj", weights, False)
            L.up = TensorParallelColumnLinear.load(c, f"{p}.mlp.up_proj", weights, False)
            L.down = TensorParallelRowLinear.load(c, f"{p}.mlp.down_proj", weights, False)
            self.layers.append(L)
        self.norm = _RMSNorm(weights.get_tensor("model.norm.weight"), c.rms_norm_eps)
        self.head = TensorParallelHead.load(c, "lm_head", weights)

    def forward(self, ids, past_key_values=None, use_cache=True):
        B, S = ids.shape
        past_len = 0 if past_key_values is None else past_key_values[0][0].shape[-2]
        pos = torch.arange(past_len, past_len + S, device=ids.device)
        x = self.embed(ids)
        presents = []
        for i, L in enumerate(self.layers):
            a, pr = L.attn(L.ln1(x), pos, None if past_key_values is None else past_key_values[i])
            presents.append(pr)
            x = x + a
            h = L.ln2(x)
            x = x + L.down(F.silu(L.gate(h)) * L.up(h))
        logits = self.head(self.norm(x))                      # all S positions + list all-gather, like the reference
        return _Out(logits, tuple(presents))


class R2Opt(nn.Module):
    def __init__(self, config, weights):
        super().__init__()
        from llmss.server.models.utils.layers import (TensorParallelColumnLinear, TensorParallelEmbedding,
                                                      TensorParallelHead, TensorParallelRowLinear)

        c = config
        d = c.hidden_size // c.num_attention_heads
        pre = "model.decoder"
        self.embed = TensorParallelEmbedding(f"{pre}.embed_tokens", weights)
        self.pos = TensorParallelEmbedding(f"{pre}.embed_positions", weights)   # learned positions: second all-reduce
        self.layers = nn.ModuleList()
        for i in range(c.num_hidden_layers):
            p = f"{pre}.layers.{i}"
            L = nn.Module()
            L.ln1 = nn.LayerNorm.load(prefix=f"{p}.self_attn_layer_norm", weights=weights, eps=1e-5)
            L.ln2 = nn.LayerNorm.load(prefix=f"{p}.final_layer_norm", weights=weights, eps=1e-5)
            L.attn = _Attn(c, f"{p}.self_attn", weights, ("q_proj", "k_proj", "v_proj", "out_proj"), True,
                           c.num_attention_heads, c.num_attention_heads, d, None)
            L.fc1 = TensorParallelColumnLinear.load(c, f"{p}.fc1", weights, True)
            L.fc2 = TensorParallelRowLinear.load(c, f"{p}.fc2", weights, True)
            self.layers.append(L)
        self.norm = nn.LayerNorm.load(prefix=f"{pre}.final_layer_norm", weights=weights, eps=1e-5)
        # tied head: the checkpoint has no lm_head tensor; the reference's head loader reads "<prefix>.weight"
        self.head = TensorParallelHead.load(c, f"{pre}.embed_tokens", weights)

    def forward(self, ids, past_key_values=None, use_cache=True):
        B, S = ids.shape
        past_len = 0 if past_key_values is None else past_key_values[0][0].shape[-2]
        pos = torch.arange(past_len, past_len + S, device=ids.device)
        x = self.embed(ids) + self.pos((pos + 2)[None, :])
        presents = []
        for i, L in enumerate(self.layers):
            a, pr = L.attn(L.ln1(x), pos, None if past_key_values is None else past_key_values[i])
            presents.append(pr)
            x = x + a
            x = x + L.fc2(F.relu(L.fc1(L.ln2(x))))
        return _Out(self.head(self.norm(x)), tuple(presents))


class _Out:
    def __init__(self, logits, past_key_values):
        self.logits, self.past_key_values = logits, past_key_values


R2_REGISTRY = {"llama": R2Llama, "opt": R2Opt}


def build_r2_model(path: str, local_rank: int, dtype=torch.float16):
    """Reference `Weights` + reference TP layers for a model family the reference has no model file for."""
    from transformers import AutoConfig

    from .ref_runner import import_reference, init_reference_dist

    _, _, weight_files, Weights = import_reference()
    process_group, rank, world = init_reference_dist()
    device = torch.device(f"cuda:{local_rank}") if torch.cuda.is_available() else torch.device("cpu")
    config = AutoConfig.from_pretrained(path)
    if config.model_type not in R2_REGISTRY:
        raise KeyError(f"no R2 assembly for model_type {config.model_type!r} (have {sorted(R2_REGISTRY)})")
    weights = Weights(weight_files(path), device=device, dtype=dtype, process_group=process_group)
    with torch.no_grad():
        model = R2_REGISTRY[config.model_type](config, weights).eval()
    return model, process_group, config
