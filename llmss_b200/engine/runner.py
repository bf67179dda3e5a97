"""ModelRunner: executes packed prefill batches and CUDA-graph-captured decode steps.

Design points vs. the reference loop (generate.py:102-145, consumer_server.py:123-166):
  * the KV cache is pre-allocated and paged; nothing is concatenated or re-allocated per token;
  * a decode step is ONE CUDA graph replay (all layers, fused collectives, LM head, sampler);
    the reference launches ~20 kernels + 2 NCCL ops per layer from Python every token;
  * every rank samples identically on-device (same logits, same counter RNG), so there is no
    per-token broadcast from rank 0 (reference N6/N10) and no device sync per token (Q13);
  * the per-step host inputs live in ONE pinned staging buffer -> one H2D copy per step, and the
    sampled tokens come back with one D2H copy.
"""
from __future__ import annotations

import os

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .. import ops
from ..models.decoder import DecoderLM
from ..utils.metrics import METRICS
from .batch import BatchInput
from .kv_cache import BlockAllocator, KVCache

_BUCKETS = (1, 2, 4, 8, 16, 24, 32, 48, 64, 96, 128, 192, 256)


@dataclass
class SamplingParams:
    """Per-request sampling controls; same fields and defaults as the reference CLI
    (generate.py:25-31) and HTTP schema (producer_server.py:9-15)."""
    max_new_tokens: int = 20
    is_greedy: bool = False
    temperature: float = 1.0
    top_p: float = 0.95
    top_k: int = 50
    seed: Optional[int] = None   # None: derived from the request id (distinct streams per request, same in any slot)

    def validate(self):
        assert self.max_new_tokens > 0, "Value of max_new_tokens should be over than 0."
        assert 0.0 < self.temperature <= 1.0, "Value of temperature is not valid."
        assert 0.0 < self.top_p <= 1.0, "Value of top_p is not valid."
        assert self.top_k >= 0, "Value of top_k is not valid."
        return self


def request_seed(params: "SamplingParams", seq_id: int) -> int:
    """31-bit RNG key of a request: the caller's seed if given, else a mix of the request id.  The sampler keys its
    counter RNG on (this, number of tokens sampled so far) and on nothing else, so a request draws the same
    tokens whichever batch slot it sits in and whoever shares the batch."""
    if params.seed is not None:
        return int(params.seed) & 0x7FFFFFFF
    x = (int(seq_id) + 1) * 0x9E3779B1 & 0xFFFFFFFF
    x ^= x >> 16
    x = x * 0x85EBCA6B & 0xFFFFFFFF
    x ^= x >> 13
    return x & 0x7FFFFFFF


class SeqState:
    __slots__ = ("seq_id", "prompt", "output", "blocks", "params", "length", "slot", "finished", "eos", "reserved",
                 "dropped")

    def __init__(self, seq_id, prompt, params, eos=None):
        self.seq_id, self.prompt, self.params = seq_id, list(prompt), params
        self.output: List[int] = []
        self.blocks: List[int] = []
        self.length = 0          # tokens whose KV is in the cache
        self.slot = -1           # row in the decode batch
        self.finished = False
        self.eos = eos
        self.reserved = 0        # cache blocks promised to this sequence at admission (worst case)
        self.dropped = 0         # sliding window: tokens (whole blocks) already evicted from the front of the cache


class ModelRunner:
    def __init__(self, core: DecoderLM, max_batch: int = 64, max_seq_len: Optional[int] = None,
                 block_size: Optional[int] = None, num_blocks: Optional[int] = None,
                 use_cuda_graph: Optional[bool] = None, prefill_chunk: Optional[int] = None,
                 sliding_window: bool = False, prefix_cache: bool = False):
        self.core, self.spec = core, core.spec
        self.device, self.dtype = core.device, core.dtype
        self.max_batch = max_batch
        self.prefill_chunk = prefill_chunk   # tokens per sequence and prefill round (None: whole prompts at once)
        self.max_seq_len = min(max_seq_len or self.spec.max_pos, self.spec.max_pos)
        # Sliding window (the working form of what the reference attempts at generate.py:133-144, quirk Q7): a sequence
        # that reaches max_seq_len keeps generating over its most recent tokens; the cache evicts whole blocks from the
        # front (block table shifted left, block returned to the pool) and positions stay absolute, which rotary
        # embeddings allow (keys keep the rotation of the position they were written at).  Families with a learned
        # position table cannot index past it, so they keep stopping at max_seq_len.
        self.sliding = bool(sliding_window) and not self.spec.learned_pos
        if block_size is None:
            # 64-token pages = one key tile of the tcgen05 decode-attention kernel (one TMA op per K / V
            # tile); the CPU oracle path and head sizes the kernel does not cover keep 16.
            block_size = 64 if (self.device.type == "cuda" and core.d in (64, 128, 256)) else 16
        self.block_size = block_size
        self.max_blocks_per_seq = -(-self.max_seq_len // block_size)
        if num_blocks is None:
            num_blocks = 1 + max_batch * self.max_blocks_per_seq
        self.kv = KVCache(self.spec.n_layers, num_blocks, block_size, core.hkv, core.d, self.dtype, self.device)
        self.alloc = BlockAllocator(num_blocks)
        self._outstanding = 0   # blocks reserved at admission and not yet allocated
        # Prefix cache: full prompt blocks stay registered after their sequence has gone; a later prompt that starts with
        # the same tokens re-uses the blocks (reference-counted) and only computes what follows -- the queries then run
        # over prior context, the paged mode of the prefill attention kernel.  The index is a trie over blocks:
        # (parent node, the block's tokens) -> (node id, block); node ids are never re-used, so entries below an
        # evicted block can no longer be reached.
        self.prefix_enabled = bool(prefix_cache)
        self._nodes: Dict[tuple, tuple] = {}
        self._block_key: Dict[int, tuple] = {}
        self._next_node = 1
        self.alloc.on_evict = self._forget_block
        self.cuda = self.device.type == "cuda"
        self.use_graph = self.cuda if use_cuda_graph is None else (use_cuda_graph and self.cuda)
        self.graphs: Dict[tuple, torch.cuda.CUDAGraph] = {}   # (padded batch, vocab-parallel greedy) -> graph
        self.launches_per_graph: Dict[int, int] = {}   # own-kernel launches captured per decode graph
        self._all_greedy = True
        self._vp_sampling_ok = True
        self._vp_plain_rows = False
        self._tuned_splits: Dict[int, int] = {}
        self.tuning_log: Dict[int, dict] = {}
        self.graph_pool = None
        self._setup_static(max_batch)
        self.stats = {"prefill_tokens": 0, "decode_steps": 0, "graph_replays": 0, "h2d_bytes": 0, "d2h_bytes": 0}

    # ------------------------------------------------------------------ static step buffers
    def _setup_static(self, B: int):
        mb = self.max_blocks_per_seq
        # ONE int32 staging buffer per direction: [tokens | positions | slots | seq_lens | block_tables].
        # The token region doubles as the sampler output, so a step is one H2D and one D2H copy.
        self.n_meta = 4 * B + B * mb
        pin = self.cuda
        self.h_in = torch.zeros(self.n_meta, dtype=torch.int32, pin_memory=pin)
        self.d_in = torch.zeros(self.n_meta, dtype=torch.int32, device=self.device)
        self.h_out = torch.zeros(B, dtype=torch.int32, pin_memory=pin)
        # The DMA engine reads a pinned source when the copy EXECUTES, not when it is enqueued, and enqueue-only
        # steps (sync=False) run many steps ahead of the GPU: each step therefore snapshots h_in into the next
        # buffer of a small ring and copies from there; a buffer is rewritten only after the event recorded behind
        # its previous copy has completed.
        self._ring = [torch.zeros(self.n_meta, dtype=torch.int32, pin_memory=pin) for _ in range(4)] if pin else []
        self._ring_ev = [None] * len(self._ring)
        self._ring_i = 0
        self.np_in = self.h_in.numpy()
        self.np_out = self.h_out.numpy()
        self.np_tok, self.np_pos = self.np_in[0:B], self.np_in[B:2 * B]
        self.np_slot, self.np_slen = self.np_in[2 * B:3 * B], self.np_in[3 * B:4 * B]
        self.np_bt = self.np_in[4 * B:].reshape(B, mb)
        self.np_len = np.zeros(B, dtype=np.int64)       # tokens already in the cache, per row
        self.np_nblk = np.zeros(B, dtype=np.int64)      # blocks owned, per row
        self.np_active = np.zeros(B, dtype=bool)
        self.np_drop = np.zeros(B, dtype=np.int64)      # sliding window: tokens evicted from the front, per row
        self.np_slot[:] = -1
        self.d_in.copy_(self.h_in)
        # sampler parameters (change only when the batch composition changes)
        self.h_samp_i = torch.zeros(2 * B, dtype=torch.int32, pin_memory=pin)     # [greedy | top_k]
        self.h_samp_f = torch.ones(2 * B, dtype=torch.float32, pin_memory=pin)    # [temperature | top_p]
        self.h_samp_i[:B] = 1
        self.d_samp_i = self.h_samp_i.to(self.device)
        self.d_samp_f = self.h_samp_f.to(self.device)
        self.d_rng = torch.zeros(B, 2, dtype=torch.int32, device=self.device)
        self.d_last_idx = torch.arange(B, dtype=torch.int64, device=self.device)
        self.d_cu = torch.arange(B + 1, dtype=torch.int32, device=self.device)
        self.batch_seqs: List[SeqState] = []

    def _dev_views(self, B: int):
        buf, MB, mb = self.d_in, self.max_batch, self.max_blocks_per_seq
        return (buf[0:B], buf[MB:MB + B], buf[2 * MB:2 * MB + B], buf[3 * MB:3 * MB + B],
                buf[4 * MB:].view(MB, mb)[:B])

    # ------------------------------------------------------------------ sequences
    def blocks_needed(self, n_tokens: int) -> int:
        return -(-n_tokens // self.block_size)

    def worst_case_blocks(self, prompt_len: int, max_new: int) -> int:
        return self.blocks_needed(min(prompt_len + max_new, self.max_seq_len))

    def fits_at_all(self, prompt_len: int, max_new: int) -> bool:
        """Could this request run alone in an empty cache?  (If not it must be rejected, not queued.)"""
        return self.worst_case_blocks(prompt_len, max_new) <= self.alloc.num_blocks - 1

    def can_admit(self, prompt_len: int, max_new: int) -> bool:
        """Admission reserves the worst case (prompt + max_new tokens): blocks promised to running sequences but
        not yet allocated are subtracted, so a decode step can never run out of cache pages half-way."""
        return self.alloc.num_free - self._outstanding >= self.worst_case_blocks(prompt_len, max_new)

    def reserve(self, seq: SeqState):
        seq.reserved = self.worst_case_blocks(min(len(seq.prompt), self.max_seq_len - 1), seq.params.max_new_tokens)
        self._outstanding += seq.reserved

    def _grow(self, seq: SeqState, total_len: int):
        need = self.blocks_needed(total_len) - len(seq.blocks)
        if need > 0:
            seq.blocks.extend(self.alloc.alloc(need))
            took = min(need, seq.reserved)
            seq.reserved -= took
            self._outstanding -= took

    def _forget_block(self, block: int):
        key = self._block_key.pop(block, None)
        if key is not None:
            self._nodes.pop(key, None)

    def _attach_cached_prefix(self, seq: SeqState) -> int:
        """Point the head of `seq`'s block table at cached blocks holding the same prompt prefix; returns the number of
        prompt tokens covered (a multiple of the block size; at least one prompt token is always left to compute,
        its logits are needed)."""
        bs, parent, got = self.block_size, 0, 0
        for i in range((len(seq.prompt) - 1) // bs):
            hit = self._nodes.get((parent, tuple(seq.prompt[i * bs:(i + 1) * bs])))
            if hit is None:
                break
            parent, blk = hit
            self.alloc.share(blk)
            seq.blocks.append(blk)
            got += bs
        took = min(got // bs, seq.reserved)      # these blocks will not be allocated after all
        seq.reserved -= took
        self._outstanding -= took
        self.stats["prefix_hit_tokens"] = self.stats.get("prefix_hit_tokens", 0) + got
        if got:
            METRICS.inc("prefix_hit_tokens", got)
        return got

    def _register_blocks(self, seq: SeqState, tokens: Sequence[int]):
        """Enter the full blocks of `tokens` (whose keys / values are in `seq`'s first blocks) into the prefix index."""
        bs, parent = self.block_size, 0
        for i in range(min(len(tokens) // bs, len(seq.blocks))):
            key = (parent, tuple(tokens[i * bs:(i + 1) * bs]))
            hit = self._nodes.get(key)
            if hit is None:
                blk = seq.blocks[i]
                if blk in self._block_key:        # already registered under another chain: cannot happen, be safe
                    break
                hit = (self._next_node, blk)
                self._next_node += 1
                self._nodes[key] = hit
                self._block_key[blk] = key
                self.alloc.keep(blk)
            parent = hit[0]

    def release(self, seq: SeqState):
        if self.prefix_enabled and seq.blocks and seq.dropped == 0 and seq.length > len(seq.prompt):
            # multi-turn reuse: what was generated is cached too (the last token's keys are not written until it is fed)
            self._register_blocks(seq, (seq.prompt + seq.output)[:seq.length])
        self.alloc.release(seq.blocks)
        seq.blocks = []
        self._outstanding -= seq.reserved
        seq.reserved = 0

    # ------------------------------------------------------------------ prefill
    @torch.no_grad()
    def prefill(self, seqs: Sequence[SeqState], rng_steps: Optional[Sequence[int]] = None) -> List[int]:
        """Run the prompts of ``seqs`` (packed, no padding) and sample their first tokens.

        With ``prefill_chunk`` set, prompts are fed in rounds of at most that many tokens per sequence: round r
        attends over the r * chunk tokens already in the cache (multi-token queries over prior context: the paged
        mode of the prefill attention kernel), which bounds the activation memory of very long prompts."""
        have = []   # prompt tokens already in the cache per sequence (prefix cache hits; else 0)
        for s in seqs:
            s.prompt = s.prompt[-(self.max_seq_len - 1):]  # left-truncate like the reference tokenizer call
            have.append(self._attach_cached_prefix(s) if self.prefix_enabled and not s.blocks else 0)
            self._grow(s, len(s.prompt))
        todo = [len(s.prompt) - h for s, h in zip(seqs, have)]
        chunk = self.prefill_chunk or max(todo)
        rounds = -(-max(todo) // chunk)
        final_logits = [None] * len(seqs)
        for r in range(rounds):
            live = [(i, s) for i, s in enumerate(seqs) if todo[i] > r * chunk]
            batch = BatchInput.build([s.prompt[have[i] + r * chunk:have[i] + (r + 1) * chunk] for i, s in live],
                                     [have[i] + r * chunk for i, _ in live],
                                     [s.blocks for _, s in live], self.block_size, self.device,
                                     max_blocks=self.max_blocks_per_seq,
                                     pad_to=self.core.tp * 16 if self.core.tp > 1 else 1)
            logits = self.core.forward(batch, self.kv, last_only=True)
            for row, (i, s) in enumerate(live):
                if todo[i] <= (r + 1) * chunk:
                    final_logits[i] = logits[row]
            self.stats["prefill_tokens"] += int(batch.tokens.numel())
        if self.prefix_enabled:
            for s in seqs:
                self._register_blocks(s, s.prompt)
        logits = torch.stack(final_logits)
        dev = self.device
        greedy = torch.tensor([int(s.params.is_greedy) for s in seqs], dtype=torch.int32, device=dev)
        temp = torch.tensor([s.params.temperature for s in seqs], dtype=torch.float32, device=dev)
        top_k = torch.tensor([s.params.top_k for s in seqs], dtype=torch.int32, device=dev)
        top_p = torch.tensor([s.params.top_p for s in seqs], dtype=torch.float32, device=dev)
        steps = list(rng_steps) if rng_steps is not None else [0] * len(seqs)
        rng = torch.tensor([[request_seed(s.params, s.seq_id), st] for s, st in zip(seqs, steps)], dtype=torch.int32, device=dev)
        toks = ops.sample(logits, greedy, temp, top_k, top_p, rng, vocab=self.spec.vocab).tolist()
        for s, t in zip(seqs, toks):
            s.length = len(s.prompt)
            s.output.append(int(t))
        return [int(t) for t in toks]

    # ------------------------------------------------------------------ decode
    def _bucket(self, B: int) -> int:
        for b in _BUCKETS:
            if b >= B and b <= self.max_batch:
                return b
        return self.max_batch

    def set_batch(self, seqs: Sequence[SeqState]):
        """(Re)bind the decode batch rows to sequences; uploads sampler parameters and block tables."""
        B, MB = len(seqs), self.max_batch
        assert B <= MB
        self.batch_seqs = list(seqs)
        self.h_samp_i[:MB] = 1
        rng = torch.zeros(MB, 2, dtype=torch.int32)
        self.np_in[:] = 0
        self.np_slot[:] = -1
        self.np_active[:] = False
        self.np_len[:] = 0
        self.np_nblk[:] = 0
        self.np_drop[:] = 0
        self._all_greedy = all(s.params.is_greedy for s in seqs)
        # vocab-parallel sampling keeps 128 candidates per rank: exact for 1 <= top_k <= 128 (ops/cuda_impl.VP_KS)
        plain = [(not s.params.is_greedy) and s.params.top_k <= 0 and s.params.top_p >= 1.0 for s in seqs]
        self._vp_plain_rows = any(plain)   # whole-distribution rows: owner pick + 16-byte gather on top of the records
        self._vp_sampling_ok = self.core.tp * 128 <= 1024 and all(
            s.params.is_greedy or 1 <= s.params.top_k <= 128 or p for s, p in zip(seqs, plain))
        for i, s in enumerate(seqs):
            s.slot = i
            self.h_samp_i[i] = int(s.params.is_greedy)
            self.h_samp_i[MB + i] = s.params.top_k
            self.h_samp_f[i] = s.params.temperature
            self.h_samp_f[MB + i] = s.params.top_p
            rng[i, 0] = request_seed(s.params, s.seq_id)
            rng[i, 1] = len(s.output)
            self.np_active[i] = not s.finished
            self.np_len[i] = s.length
            self.np_drop[i] = s.dropped
            self.np_nblk[i] = len(s.blocks)
            self.np_bt[i, :len(s.blocks)] = s.blocks
            if s.output:
                self.np_tok[i] = s.output[-1]
        self.d_samp_i.copy_(self.h_samp_i, non_blocking=True)
        self.d_samp_f.copy_(self.h_samp_f, non_blocking=True)
        self.d_rng.copy_(rng.to(self.device))

    def deactivate(self, seq: SeqState):
        """Stop generating for a bound row (its slot becomes padding until the next set_batch)."""
        seq.finished = True
        if 0 <= seq.slot < self.max_batch and seq.slot < len(self.batch_seqs) and self.batch_seqs[seq.slot] is seq:
            self.np_active[seq.slot] = False

    def _fill_step_inputs(self):
        """Vectorised (numpy on the pinned buffer) metadata for this step of the bound batch."""
        bs = self.block_size
        act = self.np_active
        pos = self.np_len                      # absolute position of the token fed this step (= tokens seen so far)
        if self.sliding:
            # window full: evict the oldest block of the row (block-granular, so slot offsets inside blocks stay put)
            full = act & (pos - self.np_drop + 1 > self.max_seq_len)
            for i in np.nonzero(full)[0]:
                s = self.batch_seqs[i]
                self.alloc.release([s.blocks.pop(0)])
                METRICS.inc("sliding_window_evicted_blocks")
                s.dropped += bs
                self.np_drop[i] += bs
                n = len(s.blocks)
                self.np_bt[i, :n] = s.blocks
                self.np_bt[i, n:] = 0
                self.np_nblk[i] = n
        w = pos - self.np_drop                 # index inside the cached window
        # rows that step onto a new cache block get one (every block_size steps per sequence)
        need = act & (w // bs >= self.np_nblk)
        for i in np.nonzero(need)[0]:
            s = self.batch_seqs[i]
            self._grow(s, int(w[i]) + 1)
            self.np_bt[i, :len(s.blocks)] = s.blocks
            self.np_nblk[i] = len(s.blocks)
        rows = np.arange(self.max_batch)
        blk = self.np_bt[rows, np.minimum(w // bs, self.max_blocks_per_seq - 1)]
        self.np_pos[:] = np.where(act, pos, 0)
        self.np_slot[:] = np.where(act, blk * bs + w % bs, -1)
        self.np_slen[:] = np.where(act, w + 1, 0)

    def _vp_greedy(self) -> int:
        """Vocab-parallel token choice (no [B, V] logits all-gather) at TP > 1 on CUDA: 1 = every bound request is
        greedy (16-byte argmax candidates), 2 = mixed greedy / top-k sampling (1 KiB candidate records), 3 = as 2 plus rows
        drawn from the whole distribution (top_k = 0, top_p = 1), 0 = a request is outside those (top_k > 128, or top_k = 0
        with top_p < 1) or the path is switched off (LLMSS_VP_GREEDY=0)."""
        c = self.core.comm
        if not (self.cuda and not c.single and c.symm is not None and os.environ.get("LLMSS_VP_GREEDY", "1") == "1"):
            return 0
        if self._all_greedy:
            return 1
        if not self._vp_sampling_ok:
            return 0
        return 3 if self._vp_plain_rows else 2

    def _decode_body(self, Bp: int, splits: int, vp_greedy: int = 0):
        tok, pos, slot, slen, bt = self._dev_views(Bp)
        MB = self.max_batch
        batch = BatchInput(tokens=tok, positions=pos, slot_mapping=slot, block_tables=bt, seq_lens=slen,
                           cu_q=self.d_cu[:Bp + 1], last_idx=self.d_last_idx[:Bp], num_seqs=Bp, max_q=1,
                           max_kv=self.max_seq_len, is_decode=True, is_fresh_prefill=False, decode_splits=splits)
        # the sampler writes into the token region: it is both this step's result and the next
        # step's input (device-side feedback loop; no host round trip needed between steps)
        if vp_greedy:
            from ..ops import cuda_impl as cu

            local = self.core.forward(batch, self.kv, last_only=False, gather_logits=False)
            if vp_greedy == 1:
                cu.greedy_vocab_parallel(local, self.core.comm.symm, self.spec.vocab, out=tok)
            else:
                cu.sample_vocab_parallel(local, self.core.comm.symm, self.spec.vocab, self.d_samp_i[:Bp],
                                         self.d_samp_f[:Bp], self.d_samp_i[MB:MB + Bp], self.d_samp_f[MB:MB + Bp],
                                         self.d_rng[:Bp], out=tok, whole_distribution_rows=(vp_greedy == 3))
            return
        logits = self.core.forward(batch, self.kv, last_only=False)  # every row is a last token
        ops.sample(logits, self.d_samp_i[:Bp], self.d_samp_f[:Bp], self.d_samp_i[MB:MB + Bp],
                   self.d_samp_f[MB:MB + Bp], self.d_rng[:Bp], vocab=self.spec.vocab, out=tok)

    def _splits_for(self, Bp: int) -> int:
        if not self.cuda:
            return 1
        if Bp in self._tuned_splits:
            return self._tuned_splits[Bp]
        from ..ops import cuda_impl as cu

        return cu.decode_splits(Bp, self.core.hq, self.core.hkv, self.core.d, self.max_seq_len, self.dtype,
                                self.block_size)

    def autotune_decode_splits(self, Bp: int) -> int:
        """Time the split-KV candidates of the decode-attention kernel on this runner's real shape
        (batch bucket, heads per rank, current context length) and keep the fastest.  Read-only on the
        KV cache; ~1 ms once per captured graph."""
        c = self.core
        if not self.cuda or c.d not in (64, 128, 256):
            return 1
        from ..ops import cuda_impl as cu

        L = int(self.np_len.max()) + 32 if self.np_active.any() else self.max_seq_len // 2
        L = max(64, min(L, self.max_seq_len))
        nblk = -(-L // self.block_size)
        dev = self.device
        bt = (torch.arange(Bp * nblk, device=dev, dtype=torch.int32) % max(1, self.kv.num_blocks - 1) + 1).view(Bp, nblk)
        sl = torch.full((Bp,), L, dtype=torch.int32, device=dev)
        q = torch.randn(Bp, c.q_dim, device=dev, dtype=self.dtype)
        out = torch.empty_like(q)
        kc, vc = self.kv.k(0), self.kv.v(0)
        base = cu.decode_splits(Bp, c.hq, c.hkv, c.d, L, self.dtype, self.block_size)
        cands = sorted({s for s in (1, 2, 3, 4, 5, 6, 8, 10, 12, 16, 24, base) if s <= max(1, L // 64)})
        best, best_t = base, float("inf")
        table = {}
        for s_ in cands:
            for _ in range(2):
                cu.attn_decode(q, kc, vc, bt, sl, out, c.hq, c.hkv, c.d, c.spec.scale, s_)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(8):
                cu.attn_decode(q, kc, vc, bt, sl, out, c.hq, c.hkv, c.d, c.spec.scale, s_)
            b.record()
            torch.cuda.synchronize()
            t = a.elapsed_time(b)
            table[s_] = round(t * 1e3 / 8, 1)
            if t < best_t:
                best, best_t = s_, t
        self._tuned_splits[Bp] = best
        self.tuning_log[Bp] = {"ctx": L, "us_per_call": table, "picked": best, "heuristic": base}
        return best

    def capture(self, Bp: int, vp: bool = False):
        """Warm up and capture the decode graph for a padded batch size, on inert inputs
        (slot -1, seq_len 0: no cache writes), then restore the live step state."""
        if not self.use_graph or (Bp, vp) in self.graphs:
            return
        splits = self._tuned_splits[Bp] if Bp in self._tuned_splits else self.autotune_decode_splits(Bp)
        saved_in, saved_rng = self.d_in.clone(), self.d_rng.clone()
        inert = torch.zeros_like(self.h_in)
        MB = self.max_batch
        inert[2 * MB:3 * MB] = -1
        self.d_in.copy_(inert)
        torch.cuda.synchronize()
        self.core.comm.barrier()
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                self._decode_body(Bp, splits, vp)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.core.comm.barrier()
        from ..ops import _lib

        g = torch.cuda.CUDAGraph()
        n0 = _lib.LAUNCHES["count"]
        with torch.cuda.graph(g, pool=self.graph_pool):
            self._decode_body(Bp, splits, vp)
        self.launches_per_graph[Bp] = _lib.LAUNCHES["count"] - n0
        if self.graph_pool is None:
            self.graph_pool = g.pool()
        self.graphs[(Bp, vp)] = g
        self.d_in.copy_(saved_in)
        self.d_rng.copy_(saved_rng)
        torch.cuda.synchronize()

    def _launch_decode(self, Bp: int):
        vp = self._vp_greedy()
        if not self.use_graph:
            self._decode_body(Bp, self._splits_for(Bp), vp)
            return
        self.graphs[(Bp, vp)].replay()
        self.stats["graph_replays"] += 1

    @torch.no_grad()
    def decode_step(self, sync: bool = True, host_tokens: bool = True) -> Optional[List[int]]:
        """One token for every bound sequence.

        ``sync=True``: full host round trip (H2D of the step inputs from pinned memory, D2H of the
        sampled tokens) -- the public serving path.  ``sync=False, host_tokens=False``: enqueue
        only; tokens stay on the device and feed the next step (device-timed throughput)."""
        B = len(self.batch_seqs)
        Bp = self._bucket(B)
        MB = self.max_batch
        if self.use_graph and (Bp, self._vp_greedy()) not in self.graphs:
            self.capture(Bp, self._vp_greedy())
        self._fill_step_inputs()
        src = self._stage()
        if host_tokens:
            self.d_in.copy_(src, non_blocking=True)
            self.stats["h2d_bytes"] += self.n_meta * 4
        else:
            self.d_in[MB:].copy_(src[MB:], non_blocking=True)
            self.stats["h2d_bytes"] += (self.n_meta - MB) * 4
        self._staged()
        self._launch_decode(Bp)
        self.stats["decode_steps"] += 1
        self.np_len[self.np_active] += 1
        if not sync:
            return None
        self.h_out[:Bp].copy_(self.d_in[:Bp], non_blocking=True)
        self.stats["d2h_bytes"] += Bp * 4
        if self.cuda:
            torch.cuda.current_stream().synchronize()
        toks = self.np_out[:B].tolist()
        self.np_tok[:B] = self.np_out[:B]
        for s, t, a in zip(self.batch_seqs, toks, self.np_active[:B].tolist()):
            if a:
                s.length += 1
                s.output.append(t)
        return toks

    def _stage(self) -> torch.Tensor:
        """A pinned snapshot of this step's inputs that nobody rewrites until its H2D copy has run."""
        if not self._ring:
            return self.h_in
        i = self._ring_i
        ev = self._ring_ev[i]
        if ev is not None:
            ev.synchronize()
        self._ring[i].copy_(self.h_in)
        return self._ring[i]

    def _staged(self):
        if not self._ring:
            return
        ev = self._ring_ev[self._ring_i]
        if ev is None:
            ev = self._ring_ev[self._ring_i] = torch.cuda.Event()
        ev.record()
        self._ring_i = (self._ring_i + 1) % len(self._ring)

    def sync_lengths(self):
        """After a run of ``sync=False`` steps: bring the host-side SeqState lengths up to date."""
        for i, s in enumerate(self.batch_seqs):
            s.length = int(self.np_len[i])
