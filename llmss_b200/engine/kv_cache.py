"""Pre-allocated paged KV cache + block allocator.

Replaces the reference's per-step ``torch.cat`` cache growth (a full O(T) copy
and a fresh allocation per layer per token; gptj_modeling.py:229-233,
gpt_bigcode_modeling.py:288-289).  Layout per layer: ``[num_blocks, kv_heads_per_rank,
block_size, head_dim]`` for K and for V, sized once for the 180 GB of HBM3e.
"""
from __future__ import annotations

from collections import OrderedDict
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
    """Reference-counted free-list allocator over cache blocks (block 0 is reserved as the padding target).

    A block is held by every sequence whose block table names it (`share` adds a holder, `release` removes one).  A
    block whose last holder leaves returns to the free list -- unless it is *pinned as cached* (`keep`: a full prompt
    block registered in the prefix index), in which case it waits in an LRU of evictable blocks and is only reclaimed,
    oldest first, when the free list is empty (`on_evict(block)` tells the index)."""

    def __init__(self, num_blocks: int, reserve_first: bool = True):
        self.num_blocks = num_blocks
        first = 1 if reserve_first else 0
        self.free: List[int] = list(range(num_blocks - 1, first - 1, -1))
        self.ref = [0] * num_blocks
        self.kept = set()                         # blocks registered in a prefix index
        self.evictable: "OrderedDict[int, None]" = OrderedDict()   # kept blocks nobody holds, least recently used first
        self.on_evict = None

    @property
    def num_free(self) -> int:
        return len(self.free) + len(self.evictable)

    def alloc(self, n: int) -> List[int]:
        if n > self.num_free:
            raise MemoryError(f"KV cache exhausted: want {n} blocks, {self.num_free} free")
        out = []
        for _ in range(n):
            if self.free:
                b = self.free.pop()
            else:
                b, _ = self.evictable.popitem(last=False)
                self.kept.discard(b)
                if self.on_evict is not None:
                    self.on_evict(b)
            self.ref[b] = 1
            out.append(b)
        return out

    def share(self, block: int):
        """One more holder of a cached block (a sequence re-using a prompt prefix)."""
        self.ref[block] += 1
        self.evictable.pop(block, None)

    def keep(self, block: int):
        self.kept.add(block)

    def release(self, blocks: List[int]):
        for b in reversed(blocks):
            self.ref[b] -= 1
            assert self.ref[b] >= 0, f"block {b} released more often than held"
            if self.ref[b] == 0:
                if b in self.kept:
                    self.evictable[b] = None      # most recently used end
                else:
                    self.free.append(b)
