"""Model registry (parity: reference custom_modeling/__init__.py:4-7) + HF-style wrappers."""
from .causal_lm import (MODEL_REGISTRY, CausalLM, CausalLMOutputWithPast, GPT2LMHeadModel, GPTBigCodeForCausalLM,
                        GPTJForCausalLM, LlamaForCausalLM, OPTForCausalLM, PagedState)
from .decoder import ArchSpec, DecoderLM

__all__ = ["MODEL_REGISTRY", "CausalLM", "CausalLMOutputWithPast", "GPTJForCausalLM", "GPTBigCodeForCausalLM",
           "GPT2LMHeadModel", "OPTForCausalLM", "LlamaForCausalLM", "PagedState", "ArchSpec", "DecoderLM"]
