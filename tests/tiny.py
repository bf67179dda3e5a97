"""Tiny HF models saved as safetensors checkpoints for CPU oracle tests."""
import torch
from transformers import (GPT2Config, GPT2LMHeadModel, GPTBigCodeConfig, GPTBigCodeForCausalLM, GPTJConfig,
                          GPTJForCausalLM, LlamaConfig, LlamaForCausalLM, OPTConfig, OPTForCausalLM)


def tiny_config(family: str, hidden: int = 64):
    """hidden=64 -> head_dim 16 (CPU oracle tests); hidden=256 -> head_dim 64 (sm_100a kernels)."""
    H = hidden
    if family == "gptj":
        return GPTJConfig(vocab_size=128, n_positions=64, n_embd=H, n_layer=2, n_head=4, rotary_dim=H // 8)
    if family == "gpt_bigcode":
        return GPTBigCodeConfig(vocab_size=128, n_positions=64, n_embd=H, n_layer=2, n_head=4, multi_query=True)
    if family == "gpt2":
        return GPT2Config(vocab_size=131, n_positions=64, n_embd=H, n_layer=2, n_head=4)
    if family == "opt":
        return OPTConfig(vocab_size=128, hidden_size=H, num_hidden_layers=2, ffn_dim=4 * H, num_attention_heads=4,
                         max_position_embeddings=64, word_embed_proj_dim=H)
    if family == "llama":
        return LlamaConfig(vocab_size=128, hidden_size=H, intermediate_size=2 * H, num_hidden_layers=2,
                           num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=64)
    raise KeyError(family)


HF_CLASSES = {"gptj": GPTJForCausalLM, "gpt_bigcode": GPTBigCodeForCausalLM, "gpt2": GPT2LMHeadModel,
              "opt": OPTForCausalLM, "llama": LlamaForCausalLM}


def make_tiny(family: str, path, seed: int = 0, hidden: int = 64):
    torch.manual_seed(seed)
    cfg = tiny_config(family, hidden)
    if family == "gpt_bigcode":
        cfg._attn_implementation = "eager"
    model = HF_CLASSES[family](cfg).eval().float()
    with torch.no_grad():  # make biases / norms non-trivial
        for n, p in model.named_parameters():
            if p.dim() == 1:
                p.add_(0.1 * torch.randn_like(p))
    model.save_pretrained(str(path), safe_serialization=True)
    return cfg, model
