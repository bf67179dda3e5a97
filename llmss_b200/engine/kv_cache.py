"""Pre-allocated paged KV cache + block allocator.

Replaces the reference's per-step ``torch.cat`` cache growth (a full O(T) copy
and a fresh allocation per layer per token; gptj_modeling.py:229-233,
gpt_bigcode_modeling.py:288-289).  Layout per layer: ``[num_blocks, kv_heads_per_rank,
block_size, head_dim]`` for K and for V, sized once for the 180 GB of HBM3e.
"""
from __future__ import annotations

from typing import List

import torch


class KVCache:
    def __init__(self, num_layers: int, num_blocks: int, block_size: int, kv_heads: int, head_dim: int, dtype, device):
        self.num_layers, self.num_blocks, self.block_size = num_layers, num_blocks, block_size
        self.kv_heads, self.head_dim = kv_heads, head_dim
        self.data = torch.zeros(num_layers, 2, num_blocks, kv_heads, block_size, head_dim, dtype=dtype, device=device)

    def k(self, layer: int) -> torch.Tensor:
        return self.data[layer, 0]

    def v(self, layer: int) -> torch.Tensor:
        return self.data[layer, 1]

    @staticmethod
    def bytes_per_block(num_layers, block_size, kv_heads, head_dim, dtype) -> int:
        return num_layers * 2 * block_size * kv_heads * head_dim * torch.empty((), dtype=dtype).element_size()


class BlockAllocator:
    """Free-list allocator over cache blocks (block 0 is reserved as the padding target)."""

    def __init__(self, num_blocks: int, reserve_first: bool = True):
        self.num_blocks = num_blocks
        first = 1 if reserve_first else 0
        self.free: List[int] = list(range(num_blocks - 1, first - 1, -1))

    @property
    def num_free(self) -> int:
        return len(self.free)

    def alloc(self, n: int) -> List[int]:
        if n > len(self.free):
            raise MemoryError(f"KV cache exhausted: want {n} blocks, {len(self.free)} free")
        return [self.free.pop() for _ in range(n)]

    def release(self, blocks: List[int]):
        self.free.extend(reversed(blocks))
