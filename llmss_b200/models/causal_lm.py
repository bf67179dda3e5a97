"""HF-shaped entry points: ``MODEL_REGISTRY[model_type](config, weights)``.

Parity with the reference's constructor / forward contract (gptj_modeling.py:517-634,
gpt_bigcode_modeling.py:786-915): ``model(input_ids, past_key_values=None, use_cache=None)``
returns an object with ``.logits`` ([B, S, V] fp32) and ``.past_key_values``.  The cache
object is an opaque :class:`PagedState` (pre-allocated paged KV, no per-step torch.cat),
which callers simply pass back, exactly as generate.py:104-106 does with HF tuples.
The serving engine does not use this wrapper; it drives ``.core`` (DecoderLM) directly.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import torch
from torch import nn

from ..engine.batch import BatchInput
from ..engine.kv_cache import KVCache
from .decoder import DecoderLM
from .families import FAMILIES


class PagedState:
    """KV state of a fixed set of B sequences advancing in lock-step (HF-compat path)."""

    def __init__(self, core: DecoderLM, batch_size: int, capacity: int, block_size: int = 16):
        self.core, self.B, self.block_size = core, batch_size, block_size
        self.length = 0
        self._alloc(capacity)

    def _alloc(self, capacity: int):
        bs = self.block_size
        self.blocks_per_seq = max(1, -(-capacity // bs))
        self.capacity = self.blocks_per_seq * bs
        c, s = self.core, self.core.spec
        self.kv = KVCache(s.n_layers, 1 + self.B * self.blocks_per_seq, bs, c.hkv, c.d, c.dtype, c.device)
        self.tables = [[1 + i * self.blocks_per_seq + j for j in range(self.blocks_per_seq)] for i in range(self.B)]

    def ensure(self, total_len: int):
        if total_len <= self.capacity:
            return
        old_kv, old_bps = self.kv, self.blocks_per_seq
        self._alloc(max(total_len, 2 * self.capacity))
        # copy the live blocks of every sequence into the enlarged cache
        for i in range(self.B):
            src = 1 + i * old_bps
            dst = 1 + i * self.blocks_per_seq
            self.kv.data[:, :, dst:dst + old_bps] = old_kv.data[:, :, src:src + old_bps]

    def get_seq_length(self) -> int:
        return self.length


@dataclass
class CausalLMOutputWithPast:
    logits: torch.Tensor
    past_key_values: Optional[PagedState] = None
    loss: Optional[torch.Tensor] = None

    def __getitem__(self, i):
        return (self.logits, self.past_key_values)[i]


class CausalLM(nn.Module):
    family: str = ""

    def __init__(self, config, weights, fuse_parallel: bool = True):
        super().__init__()
        spec_fn, load_fn = FAMILIES[self.family]
        self.config = config
        self.core = DecoderLM(spec_fn(config), weights.process_group, weights.device, weights.dtype,
                              fuse_parallel=fuse_parallel)
        load_fn(self.core, weights)
        self.core.finalize()

    @property
    def spec(self):
        return self.core.spec

    @torch.no_grad()
    def forward(self, input_ids: torch.Tensor, past_key_values: Optional[PagedState] = None,
                use_cache: Optional[bool] = None, last_only: bool = False, **_unused) -> CausalLMOutputWithPast:
        core = self.core
        B, S = input_ids.shape
        st = past_key_values
        if st is None:
            st = PagedState(core, B, capacity=min(core.spec.max_pos, S + 64))
        assert st.B == B, "batch size changed between calls"
        ctx = st.length
        st.ensure(ctx + S)
        ids = input_ids.tolist()
        batch = BatchInput.build(ids, [ctx] * B, st.tables, st.block_size, core.device, max_blocks=st.blocks_per_seq)
        hidden = core.forward_hidden(batch, st.kv)
        if last_only:
            logits = core.logits(hidden.index_select(0, batch.last_idx))[:, :core.spec.vocab].reshape(B, 1, -1)
        else:
            logits = core.logits(hidden)[:, :core.spec.vocab].reshape(B, S, -1)
        st.length = ctx + S
        return CausalLMOutputWithPast(logits=logits.float(), past_key_values=st if use_cache else None)


class GPTJForCausalLM(CausalLM):
    family = "gptj"


class GPTBigCodeForCausalLM(CausalLM):
    family = "gpt_bigcode"


class GPT2LMHeadModel(CausalLM):
    family = "gpt2"


class OPTForCausalLM(CausalLM):
    family = "opt"


class LlamaForCausalLM(CausalLM):
    family = "llama"


MODEL_REGISTRY = {
    "gptj": GPTJForCausalLM,
    "gpt_bigcode": GPTBigCodeForCausalLM,
    "gpt2": GPT2LMHeadModel,
    "opt": OPTForCausalLM,
    "llama": LlamaForCausalLM,
}
