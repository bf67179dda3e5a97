"""Checkpoint -> fused per-rank layout, one loader per model family.

Parity families (reference MODEL_REGISTRY, custom_modeling/__init__.py:4-7):
  gptj         gptj_modeling.py:84-95,259-264,281,368-377,520-524
  gpt_bigcode  gpt_bigcode_modeling.py:122-155,316-319,340-343,562-579,793-797  (MQA)
North-star families the reference cannot load (BASELINE.json configs 1-4): gpt2, opt, llama.
Key names follow the HF state dicts.
"""
from __future__ import annotations

import torch

from ..ops import ACT_IDS, NORM_LAYER, NORM_RMS, ROPE_GPTJ, ROPE_NEOX, ROPE_NONE
from .decoder import ArchSpec, DecoderLM, LayerWeights, ShardReader, interleave_gate_up


def _act(name: str) -> int:
    if name not in ACT_IDS:
        raise ValueError(f"unsupported activation {name!r}")
    return ACT_IDS[name]


def _max_pos(config) -> int:
    # quirk Q5: the reference only reads n_positions (generate.py:60)
    for k in ("max_position_embeddings", "n_positions"):
        v = getattr(config, k, None)
        if v:
            return int(v)
    raise ValueError("config has neither max_position_embeddings nor n_positions")


# ----------------------------------------------------------------------------- GPT-J
def gptj_spec(c) -> ArchSpec:
    d = c.n_embd // c.n_head
    return ArchSpec("gptj", c.n_embd, c.n_layer, c.n_head, c.n_head, d, c.n_inner or 4 * c.n_embd, c.vocab_size,
                    _max_pos(c), norm=NORM_LAYER, eps=c.layer_norm_epsilon, rope=ROPE_GPTJ,
                    rotary_dim=c.rotary_dim or d, act=_act(c.activation_function), parallel_residual=True,
                    head_bias=True, tie_head=False)


def load_gptj(m: DecoderLM, w):
    r, s = ShardReader(w, m), m.spec
    H = s.hidden
    m.wte = r.full("transformer.wte.weight")
    for i in range(s.n_layers):
        p, L = f"transformer.h.{i}", LayerWeights()
        L.ln1_w, L.ln1_b = r.full(f"{p}.ln_1.weight"), r.full(f"{p}.ln_1.bias")
        L.wqkv = torch.cat([r.q_rows(f"{p}.attn.q_proj.weight"), r.kv_rows(f"{p}.attn.k_proj.weight"),
                            r.kv_rows(f"{p}.attn.v_proj.weight")]).contiguous()
        L.wo = r.o_cols(f"{p}.attn.out_proj.weight")
        L.wup, L.bup = r.shard_rows(f"{p}.mlp.fc_in.weight", s.ffn), r.shard_rows(f"{p}.mlp.fc_in.bias", s.ffn)
        L.wdown, L.bdown = r.shard_cols(f"{p}.mlp.fc_out.weight", s.ffn), r.full(f"{p}.mlp.fc_out.bias")
        m.layers.append(L)
    m.lnf_w, m.lnf_b = r.full("transformer.ln_f.weight"), r.full("transformer.ln_f.bias")
    # quirk Q3: the reference drops lm_head.bias (layers.py:100-104); we load and apply it.
    m.w_head, m.b_head = r.head("lm_head.weight", "lm_head.bias")


# ----------------------------------------------------------------------------- GPT-BigCode
def bigcode_spec(c) -> ArchSpec:
    if not c.multi_query:
        # the reference's non-MQA path is itself broken (quirk Q8); HF interleaves q/k/v per head there
        raise ValueError("gpt_bigcode: only multi_query=True checkpoints are supported")
    d = c.n_embd // c.n_head
    return ArchSpec("gpt_bigcode", c.n_embd, c.n_layer, c.n_head, 1 if c.multi_query else c.n_head, d,
                    c.n_inner or 4 * c.n_embd, c.vocab_size, _max_pos(c), norm=NORM_LAYER, eps=c.layer_norm_epsilon,
                    learned_pos=True, act=_act(c.activation_function), tie_head=True)


def load_bigcode(m: DecoderLM, w):
    r, s = ShardReader(w, m), m.spec
    H, d = s.hidden, s.head_dim
    kvw = s.n_kv_heads * d
    m.wte, m.wpe = r.full("transformer.wte.weight"), r.full("transformer.wpe.weight")
    for i in range(s.n_layers):
        p, L = f"transformer.h.{i}", LayerWeights()
        L.ln1_w, L.ln1_b = r.full(f"{p}.ln_1.weight"), r.full(f"{p}.ln_1.bias")
        L.ln2_w, L.ln2_b = r.full(f"{p}.ln_2.weight"), r.full(f"{p}.ln_2.bias")
        # fused c_attn rows: [q (H) | k (kvw) | v (kvw)]; q head-sharded, kv sharded or replicated (lazy slices, Q9)
        parts_w = [r.q_rows(f"{p}.attn.c_attn.weight"), r.kv_rows(f"{p}.attn.c_attn.weight", H),
                   r.kv_rows(f"{p}.attn.c_attn.weight", H + kvw)]
        parts_b = [r.q_rows(f"{p}.attn.c_attn.bias"), r.kv_rows(f"{p}.attn.c_attn.bias", H),
                   r.kv_rows(f"{p}.attn.c_attn.bias", H + kvw)]
        L.wqkv, L.bqkv = torch.cat(parts_w).contiguous(), torch.cat(parts_b).contiguous()
        L.wo, L.bo = r.o_cols(f"{p}.attn.c_proj.weight"), r.full(f"{p}.attn.c_proj.bias")
        L.wup, L.bup = r.shard_rows(f"{p}.mlp.c_fc.weight", s.ffn), r.shard_rows(f"{p}.mlp.c_fc.bias", s.ffn)
        L.wdown, L.bdown = r.shard_cols(f"{p}.mlp.c_proj.weight", s.ffn), r.full(f"{p}.mlp.c_proj.bias")
        m.layers.append(L)
    m.lnf_w, m.lnf_b = r.full("transformer.ln_f.weight"), r.full("transformer.ln_f.bias")
    m.w_head, m.b_head = r.head("transformer.wte.weight")


# ----------------------------------------------------------------------------- GPT-2
def gpt2_spec(c) -> ArchSpec:
    if getattr(c, "scale_attn_by_inverse_layer_idx", False) or getattr(c, "reorder_and_upcast_attn", False):
        raise ValueError("gpt2: scale_attn_by_inverse_layer_idx / reorder_and_upcast_attn are not supported")
    d = c.n_embd // c.n_head
    return ArchSpec("gpt2", c.n_embd, c.n_layer, c.n_head, c.n_head, d, c.n_inner or 4 * c.n_embd, c.vocab_size,
                    _max_pos(c), norm=NORM_LAYER, eps=c.layer_norm_epsilon, learned_pos=True,
                    act=_act(c.activation_function), tie_head=True)


def load_gpt2(m: DecoderLM, w):
    r, s = ShardReader(w, m), m.spec
    H = s.hidden
    pre = "transformer." if w.has("transformer.wte.weight") else ""
    m.wte, m.wpe = r.full(f"{pre}wte.weight"), r.full(f"{pre}wpe.weight")
    for i in range(s.n_layers):
        p, L = f"{pre}h.{i}", LayerWeights()
        L.ln1_w, L.ln1_b = r.full(f"{p}.ln_1.weight"), r.full(f"{p}.ln_1.bias")
        L.ln2_w, L.ln2_b = r.full(f"{p}.ln_2.weight"), r.full(f"{p}.ln_2.bias")
        # Conv1D: stored [in, out]; c_attn out = [q | k | v], each H wide
        ws = [r.q_rows(f"{p}.attn.c_attn.weight", 0, transposed=True), r.kv_rows(f"{p}.attn.c_attn.weight", H, transposed=True),
              r.kv_rows(f"{p}.attn.c_attn.weight", 2 * H, transposed=True)]
        bs = [r.q_rows(f"{p}.attn.c_attn.bias", 0), r.kv_rows(f"{p}.attn.c_attn.bias", H), r.kv_rows(f"{p}.attn.c_attn.bias", 2 * H)]
        L.wqkv, L.bqkv = torch.cat(ws).contiguous(), torch.cat(bs).contiguous()
        L.wo, L.bo = r.o_cols(f"{p}.attn.c_proj.weight", transposed=True), r.full(f"{p}.attn.c_proj.bias")
        L.wup = r.shard_rows(f"{p}.mlp.c_fc.weight", s.ffn, transposed=True)
        L.bup = r.shard_rows(f"{p}.mlp.c_fc.bias", s.ffn)
        L.wdown, L.bdown = r.shard_cols(f"{p}.mlp.c_proj.weight", s.ffn, transposed=True), r.full(f"{p}.mlp.c_proj.bias")
        m.layers.append(L)
    m.lnf_w, m.lnf_b = r.full(f"{pre}ln_f.weight"), r.full(f"{pre}ln_f.bias")
    m.w_head, m.b_head = r.head(f"{pre}wte.weight")


# ----------------------------------------------------------------------------- OPT
def opt_spec(c) -> ArchSpec:
    if not c.do_layer_norm_before:
        raise ValueError("opt: do_layer_norm_before=False (OPT-350m) is not supported")
    if c.word_embed_proj_dim != c.hidden_size:
        raise ValueError("opt: word_embed_proj_dim != hidden_size is not supported")
    d = c.hidden_size // c.num_attention_heads
    return ArchSpec("opt", c.hidden_size, c.num_hidden_layers, c.num_attention_heads, c.num_attention_heads, d,
                    c.ffn_dim, c.vocab_size, _max_pos(c), norm=NORM_LAYER, eps=1e-5, learned_pos=True, pos_offset=2,
                    act=_act(c.activation_function), tie_head=True)


def load_opt(m: DecoderLM, w):
    r, s = ShardReader(w, m), m.spec
    H = s.hidden
    pre = "model.decoder." if w.has("model.decoder.embed_tokens.weight") else "decoder."
    m.wte, m.wpe = r.full(f"{pre}embed_tokens.weight"), r.full(f"{pre}embed_positions.weight")
    for i in range(s.n_layers):
        p, L = f"{pre}layers.{i}", LayerWeights()
        L.ln1_w, L.ln1_b = r.full(f"{p}.self_attn_layer_norm.weight"), r.full(f"{p}.self_attn_layer_norm.bias")
        L.ln2_w, L.ln2_b = r.full(f"{p}.final_layer_norm.weight"), r.full(f"{p}.final_layer_norm.bias")
        sa = f"{p}.self_attn"
        L.wqkv = torch.cat([r.q_rows(f"{sa}.q_proj.weight"), r.kv_rows(f"{sa}.k_proj.weight"),
                            r.kv_rows(f"{sa}.v_proj.weight")]).contiguous()
        if w.has(f"{sa}.q_proj.bias"):
            L.bqkv = torch.cat([r.q_rows(f"{sa}.q_proj.bias"), r.kv_rows(f"{sa}.k_proj.bias"),
                                r.kv_rows(f"{sa}.v_proj.bias")]).contiguous()
        L.wo, L.bo = r.o_cols(f"{sa}.out_proj.weight"), r.maybe(f"{sa}.out_proj.bias")
        L.wup = r.shard_rows(f"{p}.fc1.weight", s.ffn)
        L.bup = r.shard_rows(f"{p}.fc1.bias", s.ffn) if w.has(f"{p}.fc1.bias") else None
        L.wdown, L.bdown = r.shard_cols(f"{p}.fc2.weight", s.ffn), r.maybe(f"{p}.fc2.bias")
        m.layers.append(L)
    m.lnf_w, m.lnf_b = r.full(f"{pre}final_layer_norm.weight"), r.full(f"{pre}final_layer_norm.bias")
    head = "lm_head.weight" if w.has("lm_head.weight") else f"{pre}embed_tokens.weight"
    m.w_head, m.b_head = r.head(head)


# ----------------------------------------------------------------------------- Llama
def llama_spec(c) -> ArchSpec:
    d = getattr(c, "head_dim", None) or c.hidden_size // c.num_attention_heads
    theta = getattr(c, "rope_theta", None)
    if theta is None:
        rp = getattr(c, "rope_parameters", None) or {}
        theta = rp.get("rope_theta", 10000.0)
        if rp.get("rope_type", "default") not in ("default", None):
            raise ValueError(f"llama: rope_type {rp.get('rope_type')!r} is not supported")
    return ArchSpec("llama", c.hidden_size, c.num_hidden_layers, c.num_attention_heads,
                    getattr(c, "num_key_value_heads", None) or c.num_attention_heads, d, c.intermediate_size,
                    c.vocab_size, _max_pos(c), norm=NORM_RMS, eps=c.rms_norm_eps, rope=ROPE_NEOX, rotary_dim=d,
                    rope_base=float(theta), act=_act(c.hidden_act), gated=True,
                    tie_head=bool(getattr(c, "tie_word_embeddings", False)))


def load_llama(m: DecoderLM, w):
    r, s = ShardReader(w, m), m.spec
    d = s.head_dim
    m.wte = r.full("model.embed_tokens.weight")
    for i in range(s.n_layers):
        p, L = f"model.layers.{i}", LayerWeights()
        L.ln1_w, L.ln2_w = r.full(f"{p}.input_layernorm.weight"), r.full(f"{p}.post_attention_layernorm.weight")
        sa = f"{p}.self_attn"
        L.wqkv = torch.cat([r.q_rows(f"{sa}.q_proj.weight"), r.kv_rows(f"{sa}.k_proj.weight"),
                            r.kv_rows(f"{sa}.v_proj.weight")]).contiguous()
        if w.has(f"{sa}.q_proj.bias"):
            L.bqkv = torch.cat([r.q_rows(f"{sa}.q_proj.bias"), r.kv_rows(f"{sa}.k_proj.bias"),
                                r.kv_rows(f"{sa}.v_proj.bias")]).contiguous()
        L.wo, L.bo = r.o_cols(f"{sa}.o_proj.weight"), r.maybe(f"{sa}.o_proj.bias")
        L.wup = interleave_gate_up(r.shard_rows(f"{p}.mlp.gate_proj.weight", s.ffn), r.shard_rows(f"{p}.mlp.up_proj.weight", s.ffn))
        if w.has(f"{p}.mlp.gate_proj.bias"):
            L.bup = interleave_gate_up(r.shard_rows(f"{p}.mlp.gate_proj.bias", s.ffn), r.shard_rows(f"{p}.mlp.up_proj.bias", s.ffn))
        L.wdown, L.bdown = r.shard_cols(f"{p}.mlp.down_proj.weight", s.ffn), r.maybe(f"{p}.mlp.down_proj.bias")
        m.layers.append(L)
    m.lnf_w = r.full("model.norm.weight")
    head = "lm_head.weight" if (w.has("lm_head.weight") and not s.tie_head) else "model.embed_tokens.weight"
    m.w_head, m.b_head = r.head(head)


FAMILIES = {
    "gptj": (gptj_spec, load_gptj),
    "gpt_bigcode": (bigcode_spec, load_bigcode),
    "gpt2": (gpt2_spec, load_gpt2),
    "opt": (opt_spec, load_opt),
    "llama": (llama_spec, load_llama),
}
