"""Random-init HF-format checkpoints (config.json + *.safetensors) for benches and tests.

There is no network for real checkpoints, so benchmarks use random weights of the named
architecture written in the ordinary unsharded HF layout -- the same
files are then loaded by this framework *and* by the unmodified reference through its
`Weights` class.  Tensors are generated shard by shard on `device` (GPU when available) so a
6B-parameter checkpoint takes seconds, not minutes.
"""
from __future__ import annotations

import json
import os
from pathlib import Path
from typing import Dict, Iterator, List, Tuple

import torch
from safetensors.torch import save_file

PRESETS: Dict[str, dict] = {
    # GPT-J-6B shape: the largest architecture the reference itself can load
    "gptj-6b": dict(model_type="gptj", architectures=["GPTJForCausalLM"], vocab_size=50400, n_positions=2048,
                    n_embd=4096, n_layer=28, n_head=16, rotary_dim=64, n_inner=None, activation_function="gelu_new",
                    layer_norm_epsilon=1e-5, tie_word_embeddings=False, bos_token_id=50256, eos_token_id=50256),
    "gptj-tiny": dict(model_type="gptj", architectures=["GPTJForCausalLM"], vocab_size=512, n_positions=256,
                      n_embd=256, n_layer=2, n_head=4, rotary_dim=32, n_inner=None, activation_function="gelu_new",
                      layer_norm_epsilon=1e-5, tie_word_embeddings=False, bos_token_id=0, eos_token_id=0),
    "gptj-tiny8": dict(model_type="gptj", architectures=["GPTJForCausalLM"], vocab_size=512, n_positions=256,
                       n_embd=512, n_layer=2, n_head=8, rotary_dim=32, n_inner=None, activation_function="gelu_new",
                       layer_norm_epsilon=1e-5, tie_word_embeddings=False, bos_token_id=0, eos_token_id=0),
    # 8-rank test models: GQA (16 q heads over 8 kv heads: one kv head per rank at TP=8) and MQA (kv replicated)
    "llama-tiny8": dict(model_type="llama", architectures=["LlamaForCausalLM"], vocab_size=512, hidden_size=1024,
                        intermediate_size=2048, num_hidden_layers=2, num_attention_heads=16, num_key_value_heads=8,
                        max_position_embeddings=256, rms_norm_eps=1e-5, hidden_act="silu", rope_theta=10000.0,
                        tie_word_embeddings=False),
    "bigcode-tiny8": dict(model_type="gpt_bigcode", architectures=["GPTBigCodeForCausalLM"], vocab_size=512,
                          n_positions=256, n_embd=512, n_layer=2, n_head=8, n_inner=2048, multi_query=True,
                          activation_function="gelu_pytorch_tanh", layer_norm_epsilon=1e-5, tie_word_embeddings=True),
    # StarCoder shape (MQA) -- the reference's second family
    "starcoder": dict(model_type="gpt_bigcode", architectures=["GPTBigCodeForCausalLM"], vocab_size=49152,
                      n_positions=8192, n_embd=6144, n_layer=40, n_head=48, n_inner=24576, multi_query=True,
                      activation_function="gelu_pytorch_tanh", layer_norm_epsilon=1e-5, tie_word_embeddings=True),
    "llama-2-7b": dict(model_type="llama", architectures=["LlamaForCausalLM"], vocab_size=32000, hidden_size=4096,
                       intermediate_size=11008, num_hidden_layers=32, num_attention_heads=32, num_key_value_heads=32,
                       max_position_embeddings=4096, rms_norm_eps=1e-5, hidden_act="silu", rope_theta=10000.0,
                       tie_word_embeddings=False),
    "llama-2-13b": dict(model_type="llama", architectures=["LlamaForCausalLM"], vocab_size=32000, hidden_size=5120,
                        intermediate_size=13824, num_hidden_layers=40, num_attention_heads=40, num_key_value_heads=40,
                        max_position_embeddings=4096, rms_norm_eps=1e-5, hidden_act="silu", rope_theta=10000.0,
                        tie_word_embeddings=False),
    "opt-1.3b": dict(model_type="opt", architectures=["OPTForCausalLM"], vocab_size=50272, hidden_size=2048,
                     num_hidden_layers=24, ffn_dim=8192, num_attention_heads=32, max_position_embeddings=2048,
                     do_layer_norm_before=True, word_embed_proj_dim=2048, activation_function="relu",
                     tie_word_embeddings=True),
    # GPT-2 XL: 25 heads -- not divisible by 2; BASELINE config 2 names it at TP=2 (heads are padded, models/decoder.py)
    "gpt2-xl": dict(model_type="gpt2", architectures=["GPT2LMHeadModel"], vocab_size=50257, n_positions=1024, n_embd=1600,
                    n_layer=48, n_head=25, activation_function="gelu_new", layer_norm_epsilon=1e-5,
                    tie_word_embeddings=True),
    "gpt2": dict(model_type="gpt2", architectures=["GPT2LMHeadModel"], vocab_size=50257, n_positions=1024, n_embd=768,
                 n_layer=12, n_head=12, activation_function="gelu_new", layer_norm_epsilon=1e-5,
                 tie_word_embeddings=True),
}


def _shapes(cfg: dict) -> Iterator[Tuple[str, Tuple[int, ...], str]]:
    """(name, shape, kind) for every tensor of the family; kind in {w, b, ln_w, ln_b, emb}."""
    mt = cfg["model_type"]
    if mt == "gptj":
        H, L, V = cfg["n_embd"], cfg["n_layer"], cfg["vocab_size"]
        F = cfg.get("n_inner") or 4 * H
        yield "transformer.wte.weight", (V, H), "emb"
        for i in range(L):
            p = f"transformer.h.{i}"
            yield f"{p}.ln_1.weight", (H,), "ln_w"
            yield f"{p}.ln_1.bias", (H,), "ln_b"
            for n in ("q_proj", "k_proj", "v_proj", "out_proj"):
                yield f"{p}.attn.{n}.weight", (H, H), "w"
            yield f"{p}.mlp.fc_in.weight", (F, H), "w"
            yield f"{p}.mlp.fc_in.bias", (F,), "b"
            yield f"{p}.mlp.fc_out.weight", (H, F), "w"
            yield f"{p}.mlp.fc_out.bias", (H,), "b"
        yield "transformer.ln_f.weight", (H,), "ln_w"
        yield "transformer.ln_f.bias", (H,), "ln_b"
        yield "lm_head.weight", (V, H), "w"
        yield "lm_head.bias", (V,), "b"
    elif mt == "gpt_bigcode":
        H, L, V, P = cfg["n_embd"], cfg["n_layer"], cfg["vocab_size"], cfg["n_positions"]
        F = cfg.get("n_inner") or 4 * H
        d = H // cfg["n_head"]
        yield "transformer.wte.weight", (V, H), "emb"
        yield "transformer.wpe.weight", (P, H), "emb"
        for i in range(L):
            p = f"transformer.h.{i}"
            for ln in ("ln_1", "ln_2"):
                yield f"{p}.{ln}.weight", (H,), "ln_w"
                yield f"{p}.{ln}.bias", (H,), "ln_b"
            yield f"{p}.attn.c_attn.weight", (H + 2 * d, H), "w"
            yield f"{p}.attn.c_attn.bias", (H + 2 * d,), "b"
            yield f"{p}.attn.c_proj.weight", (H, H), "w"
            yield f"{p}.attn.c_proj.bias", (H,), "b"
            yield f"{p}.mlp.c_fc.weight", (F, H), "w"
            yield f"{p}.mlp.c_fc.bias", (F,), "b"
            yield f"{p}.mlp.c_proj.weight", (H, F), "w"
            yield f"{p}.mlp.c_proj.bias", (H,), "b"
        yield "transformer.ln_f.weight", (H,), "ln_w"
        yield "transformer.ln_f.bias", (H,), "ln_b"
    elif mt == "gpt2":
        H, L, V, P = cfg["n_embd"], cfg["n_layer"], cfg["vocab_size"], cfg["n_positions"]
        F = cfg.get("n_inner") or 4 * H
        yield "transformer.wte.weight", (V, H), "emb"
        yield "transformer.wpe.weight", (P, H), "emb"
        for i in range(L):
            p = f"transformer.h.{i}"
            for ln in ("ln_1", "ln_2"):
                yield f"{p}.{ln}.weight", (H,), "ln_w"
                yield f"{p}.{ln}.bias", (H,), "ln_b"
            yield f"{p}.attn.c_attn.weight", (H, 3 * H), "w"   # Conv1D: [in, out]
            yield f"{p}.attn.c_attn.bias", (3 * H,), "b"
            yield f"{p}.attn.c_proj.weight", (H, H), "w"
            yield f"{p}.attn.c_proj.bias", (H,), "b"
            yield f"{p}.mlp.c_fc.weight", (H, F), "w"
            yield f"{p}.mlp.c_fc.bias", (F,), "b"
            yield f"{p}.mlp.c_proj.weight", (F, H), "w"
            yield f"{p}.mlp.c_proj.bias", (H,), "b"
        yield "transformer.ln_f.weight", (H,), "ln_w"
        yield "transformer.ln_f.bias", (H,), "ln_b"
    elif mt == "opt":
        H, L, V, P, F = (cfg["hidden_size"], cfg["num_hidden_layers"], cfg["vocab_size"],
                         cfg["max_position_embeddings"], cfg["ffn_dim"])
        yield "model.decoder.embed_tokens.weight", (V, H), "emb"
        yield "model.decoder.embed_positions.weight", (P + 2, H), "emb"
        for i in range(L):
            p = f"model.decoder.layers.{i}"
            for n in ("q_proj", "k_proj", "v_proj", "out_proj"):
                yield f"{p}.self_attn.{n}.weight", (H, H), "w"
                yield f"{p}.self_attn.{n}.bias", (H,), "b"
            for ln in ("self_attn_layer_norm", "final_layer_norm"):
                yield f"{p}.{ln}.weight", (H,), "ln_w"
                yield f"{p}.{ln}.bias", (H,), "ln_b"
            yield f"{p}.fc1.weight", (F, H), "w"
            yield f"{p}.fc1.bias", (F,), "b"
            yield f"{p}.fc2.weight", (H, F), "w"
            yield f"{p}.fc2.bias", (H,), "b"
        yield "model.decoder.final_layer_norm.weight", (H,), "ln_w"
        yield "model.decoder.final_layer_norm.bias", (H,), "ln_b"
    elif mt == "llama":
        H, L, V, F = cfg["hidden_size"], cfg["num_hidden_layers"], cfg["vocab_size"], cfg["intermediate_size"]
        hq, hkv = cfg["num_attention_heads"], cfg.get("num_key_value_heads") or cfg["num_attention_heads"]
        d = cfg.get("head_dim") or H // hq
        yield "model.embed_tokens.weight", (V, H), "emb"
        for i in range(L):
            p = f"model.layers.{i}"
            yield f"{p}.input_layernorm.weight", (H,), "ln_w"
            yield f"{p}.post_attention_layernorm.weight", (H,), "ln_w"
            yield f"{p}.self_attn.q_proj.weight", (hq * d, H), "w"
            yield f"{p}.self_attn.k_proj.weight", (hkv * d, H), "w"
            yield f"{p}.self_attn.v_proj.weight", (hkv * d, H), "w"
            yield f"{p}.self_attn.o_proj.weight", (H, hq * d), "w"
            yield f"{p}.mlp.gate_proj.weight", (F, H), "w"
            yield f"{p}.mlp.up_proj.weight", (F, H), "w"
            yield f"{p}.mlp.down_proj.weight", (H, F), "w"
        yield "model.norm.weight", (H,), "ln_w"
        if not cfg.get("tie_word_embeddings", False):
            yield "lm_head.weight", (V, H), "w"
    else:
        raise KeyError(mt)


def checkpoint_bytes(cfg: dict, elem_bytes: int = 2) -> int:
    n = 0
    for _, shape, _ in _shapes(cfg):
        k = 1
        for s in shape:
            k *= s
        n += k
    return n * elem_bytes


def write_random_checkpoint(path, preset_or_cfg, dtype=torch.float16, device=None, seed: int = 0,
                            shard_bytes: int = 2 << 30, std: float = 0.02) -> Path:
    """Create ``path/config.json`` and ``path/model-XXXXX.safetensors``; idempotent (a marker
    file records completion so concurrent ranks / reruns can skip the work)."""
    cfg = dict(PRESETS[preset_or_cfg]) if isinstance(preset_or_cfg, str) else dict(preset_or_cfg)
    path = Path(path)
    done = path / ".complete"
    if done.exists():
        return path
    path.mkdir(parents=True, exist_ok=True)
    cfg.setdefault("torch_dtype", str(dtype).replace("torch.", ""))
    (path / "config.json").write_text(json.dumps(cfg, indent=1))
    if device is None:
        device = "cuda" if torch.cuda.is_available() else "cpu"
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    shard: Dict[str, torch.Tensor] = {}
    size, idx, total = 0, 0, 0
    weight_map: Dict[str, str] = {}

    def flush():
        nonlocal shard, size, idx
        if shard:
            fname = f"model-{idx:05d}.safetensors"
            save_file({k: v.cpu() for k, v in shard.items()}, str(path / fname), metadata={"format": "pt"})
            weight_map.update({k: fname for k in shard})
            idx += 1
            shard, size = {}, 0

    for name, shape, kind in _shapes(cfg):
        if kind == "ln_w":
            t = 1.0 + 0.05 * torch.randn(shape, generator=gen, device=device, dtype=torch.float32)
        elif kind in ("b", "ln_b"):
            t = 0.02 * torch.randn(shape, generator=gen, device=device, dtype=torch.float32)
        else:
            t = std * torch.randn(shape, generator=gen, device=device, dtype=torch.float32)
        shard[name] = t.to(dtype)
        size += shard[name].numel() * shard[name].element_size()
        total += shard[name].numel() * shard[name].element_size()
        if size >= shard_bytes:
            flush()
    flush()
    (path / "model.safetensors.index.json").write_text(
        json.dumps({"metadata": {"total_size": total}, "weight_map": weight_map}, indent=1))
    done.write_text("ok")
    return path
