"""Packed (varlen) batch description shared by models, engine and ops."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch


@dataclass
class BatchInput:
    tokens: torch.Tensor        # [T] int32 token ids, sequences packed back to back
    positions: torch.Tensor     # [T] int32 absolute position of each token
    slot_mapping: torch.Tensor  # [T] int32 KV-cache slot (block * block_size + offset); < 0 = do not write
    block_tables: torch.Tensor  # [B, max_blocks] int32
    seq_lens: torch.Tensor      # [B] int32 total cached length *including* this step's tokens
    cu_q: torch.Tensor          # [B + 1] int32 query row offsets
    last_idx: torch.Tensor      # [B] int64 row of the last token of each sequence
    num_seqs: int
    max_q: int
    max_kv: int
    is_decode: bool             # exactly one query token per sequence
    is_fresh_prefill: bool      # no sequence has prior context (seq_len == q_len)
    decode_splits: int = 0      # split-KV factor frozen at graph-capture time (0 = choose per call)

    @staticmethod
    def build(token_lists: Sequence[Sequence[int]], context_lens: Sequence[int], block_tables: Sequence[Sequence[int]],
              block_size: int, device, max_blocks: Optional[int] = None, pad_to: int = 1) -> "BatchInput":
        """Host-side construction: ``token_lists[i]`` are the new tokens of sequence i, which
        already has ``context_lens[i]`` tokens in the cache blocks ``block_tables[i]``."""
        toks: List[int] = []
        pos: List[int] = []
        slots: List[int] = []
        cu = [0]
        seq_lens = []
        for new, ctx, bt in zip(token_lists, context_lens, block_tables):
            for j, t in enumerate(new):
                p = ctx + j
                toks.append(int(t))
                pos.append(p)
                slots.append(bt[p // block_size] * block_size + p % block_size)
            cu.append(len(toks))
            seq_lens.append(ctx + len(new))
        # inert padding rows (token 0, slot -1, owned by no sequence) so that the packed length is a
        # multiple of `pad_to` (sequence-parallel prefill shards the rows evenly over the ranks)
        while len(toks) % pad_to:
            toks.append(0)
            pos.append(0)
            slots.append(-1)
        B = len(token_lists)
        mb = max_blocks or max(1, max(len(b) for b in block_tables))
        bt_t = torch.zeros(B, mb, dtype=torch.int32)
        for i, b in enumerate(block_tables):
            if len(b):
                bt_t[i, :len(b)] = torch.tensor(b, dtype=torch.int32)
        qlens = [len(t) for t in token_lists]
        i32 = dict(dtype=torch.int32)
        return BatchInput(
            tokens=torch.tensor(toks, **i32).to(device), positions=torch.tensor(pos, **i32).to(device),
            slot_mapping=torch.tensor(slots, **i32).to(device), block_tables=bt_t.to(device),
            seq_lens=torch.tensor(seq_lens, **i32).to(device), cu_q=torch.tensor(cu, **i32).to(device),
            last_idx=(torch.tensor(cu[1:], dtype=torch.int64) - 1).to(device),
            num_seqs=B, max_q=max(qlens) if qlens else 0, max_kv=max(seq_lens) if seq_lens else 0,
            is_decode=all(q == 1 for q in qlens), is_fresh_prefill=all(c == 0 for c in context_lens))
