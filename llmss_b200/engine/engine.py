"""Engine: the public serving API -- request admission, continuous batching, generation.

The reference has no engine: `generate.py` is a fixed-batch loop and the PoC consumer serves
one request at a time with batch size 1 (consumer_server.py:73).  Here requests join and
leave the running batch at step boundaries (continuous batching), prompts are prefetched as
packed varlen batches, and decode runs as CUDA-graph replays (see runner.py).

Every TP rank runs an identical Engine on identical inputs, so scheduling decisions are
replicated deterministically and need no per-step collective; only the request payload has to
be shared (generate.py broadcasts the prompt ids; the consumer broadcasts the request plan).
"""
from __future__ import annotations

import time
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import torch

from ..models import MODEL_REGISTRY
from ..parallel.comm import comm_for
from ..utils.dist import FakeGroup
from ..utils.hub import weight_files
from ..utils.metrics import METRICS
from ..utils.weights import Weights
from .runner import ModelRunner, SamplingParams, SeqState


@dataclass
class StepOutput:
    request_id: int
    token: int
    finished: bool


@dataclass
class RequestStats:
    arrival: float
    first_token: Optional[float] = None
    last_token: Optional[float] = None
    n_out: int = 0

    @property
    def ttft(self):
        return None if self.first_token is None else self.first_token - self.arrival

    @property
    def tpot(self):
        if self.n_out < 2 or self.last_token is None:
            return None
        return (self.last_token - self.first_token) / (self.n_out - 1)


class Engine:
    def __init__(self, model, max_batch: int = 64, max_seq_len: Optional[int] = None,
                 block_size: Optional[int] = None,
                 num_blocks: Optional[int] = None, use_cuda_graph: Optional[bool] = None,
                 max_prefill_tokens: int = 16384, eos_token_id: Optional[int] = None,
                 prefill_chunk: Optional[int] = None, sliding_window: bool = False, prefix_cache: bool = False):
        self.model = model
        self.core = model.core if hasattr(model, "core") else model
        self.comm = self.core.comm
        self.runner = ModelRunner(self.core, max_batch=max_batch, max_seq_len=max_seq_len, block_size=block_size,
                                  num_blocks=num_blocks, use_cuda_graph=use_cuda_graph, prefill_chunk=prefill_chunk,
                                  sliding_window=sliding_window, prefix_cache=prefix_cache)
        self.max_batch = max_batch
        self.max_prefill_tokens = max_prefill_tokens
        self.eos_token_id = eos_token_id
        self.waiting: List[SeqState] = []
        self.running: List[SeqState] = []
        self.stats: Dict[int, RequestStats] = {}
        self._next_id = 0
        self.rejected: List[int] = []   # requests that cannot fit the KV cache even alone (reported by step())
        self._dirty = True  # decode batch needs re-binding

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_pretrained(cls, path: str, dtype=torch.bfloat16, device=None, process_group=None,
                        symmetric_bytes: int = 128 << 20, lamport_bytes: int = 64 << 20, fuse_parallel: bool = True,
                        **kw) -> "Engine":
        from transformers import AutoConfig

        if process_group is None:
            process_group = FakeGroup(0, 1)
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        device = torch.device(device)
        if device.type == "cpu" and dtype in (torch.float16, torch.bfloat16):
            dtype = torch.float32  # CPU oracle path runs in fp32
        config = AutoConfig.from_pretrained(path)
        comm = comm_for(process_group, device)
        if device.type == "cuda" and comm.size > 1 and comm.symm is None:
            comm.attach_symmetric(symmetric_bytes, lamport_bytes)
        weights = Weights(weight_files(path), device=device, dtype=dtype, process_group=process_group)
        if config.model_type not in MODEL_REGISTRY:
            raise KeyError(f"model_type {config.model_type!r} is not supported; have {sorted(MODEL_REGISTRY)}")
        model = MODEL_REGISTRY[config.model_type](config, weights, fuse_parallel=fuse_parallel).eval()
        weights.close()
        return cls(model, **kw)

    # ------------------------------------------------------------------ requests
    def add_request(self, prompt_ids: Sequence[int], params: Optional[SamplingParams] = None,
                    request_id: Optional[int] = None) -> int:
        params = (params or SamplingParams()).validate()
        rid = self._next_id if request_id is None else request_id
        self._next_id = max(self._next_id, rid) + 1
        if len(prompt_ids) == 0:
            raise ValueError("empty prompt")
        self.waiting.append(SeqState(rid, prompt_ids, params, eos=self.eos_token_id))
        self.stats[rid] = RequestStats(arrival=time.perf_counter())
        METRICS.inc("requests_total")
        return rid

    def has_work(self) -> bool:
        return bool(self.waiting or self.running)

    def abort_all(self) -> List[int]:
        """Drop every waiting and running request and give their KV blocks back (failure recovery / shutdown).
        Returns the ids that were dropped; on a TP job every rank must call it at the same point."""
        dropped = [s.seq_id for s in self.waiting] + [s.seq_id for s in self.running if not s.finished]
        for s in self.running:
            if not s.finished:
                s.finished = True
                self._finish(s)
        self.waiting.clear()
        self.running = []
        self.runner.set_batch([])
        self._dirty = True
        for rid in dropped:
            self.stats.pop(rid, None)
        return dropped

    # ------------------------------------------------------------------ scheduling
    def _admit(self) -> List[SeqState]:
        admitted, budget = [], self.max_prefill_tokens
        r = self.runner
        while self.waiting and len(self.running) + len(admitted) < self.max_batch:
            s = self.waiting[0]
            n = min(len(s.prompt), r.max_seq_len - 1)
            if admitted and n > budget:
                break
            if not r.fits_at_all(n, s.params.max_new_tokens):
                # can never run, not even alone in an empty cache: reject instead of blocking the queue for ever
                self.waiting.pop(0)
                s.finished = True
                self.rejected.append(s.seq_id)
                METRICS.inc("requests_rejected")
                continue
            if not r.can_admit(n, s.params.max_new_tokens):
                break
            budget -= n
            s = self.waiting.pop(0)
            r.reserve(s)
            admitted.append(s)
        return admitted

    def _is_done(self, s: SeqState, tok: int) -> bool:
        return (len(s.output) >= s.params.max_new_tokens
                or (s.eos is not None and tok == s.eos)
                or (not self.runner.sliding and len(s.prompt) + len(s.output) >= self.runner.max_seq_len))

    def _finish(self, s: SeqState):
        self.runner.deactivate(s)
        self.runner.release(s)
        self._dirty = True
        METRICS.inc("requests_finished")

    def step(self) -> List[StepOutput]:
        """One scheduling step: prefill newly admitted requests if any, else one decode step."""
        now = time.perf_counter
        outs: List[StepOutput] = []
        new = self._admit()
        while self.rejected:  # surfaced as finished outputs with token -1 (the servers turn them into errors)
            outs.append(StepOutput(self.rejected.pop(0), -1, True))
        if new:
            t0 = now()
            toks = self.runner.prefill(new)
            t1 = now()
            METRICS.observe("prefill_seconds", t1 - t0)
            METRICS.inc("prefill_tokens", sum(len(s.prompt) for s in new))
            for s, t in zip(new, toks):
                st = self.stats[s.seq_id]
                st.first_token = st.last_token = t1
                st.n_out = 1
                METRICS.observe("ttft_seconds", st.ttft)
                done = self._is_done(s, t)
                outs.append(StepOutput(s.seq_id, t, done))
                if done:
                    s.finished = True
                    self.runner.release(s)
                    METRICS.inc("requests_finished")
                else:
                    self.running.append(s)
            self._dirty = True
            return outs
        if not self.running:
            return outs
        if self._dirty:
            self.running = [s for s in self.running if not s.finished]
            self.runner.set_batch(self.running)
            self._dirty = False
            if not self.running:
                return outs
        t0 = now()
        toks = self.runner.decode_step(sync=True)
        t1 = now()
        METRICS.observe("decode_step_seconds", t1 - t0)
        METRICS.set("batch_occupancy", len(self.running))
        a = self.runner.alloc
        METRICS.set("kv_blocks_free", a.num_free)               # free list + evictable cached blocks
        METRICS.set("kv_blocks_cached", len(a.kept))            # blocks registered in the prefix index
        METRICS.set("requests_waiting", len(self.waiting))
        n_active = 0
        for s, t in zip(self.running, toks):
            if s.finished:
                continue
            n_active += 1
            st = self.stats[s.seq_id]
            st.last_token, st.n_out = t1, st.n_out + 1
            done = self._is_done(s, t)
            outs.append(StepOutput(s.seq_id, t, done))
            if done:
                METRICS.observe("tpot_seconds", st.tpot or 0.0)
                self._finish(s)
        METRICS.inc("decode_tokens", n_active)
        return outs

    # ------------------------------------------------------------------ convenience
    def generate(self, prompts: Sequence[Sequence[int]], params=None) -> List[List[int]]:
        """Blocking batch generation: returns the new tokens of every prompt, in order."""
        if params is None or isinstance(params, SamplingParams):
            params = [params or SamplingParams()] * len(prompts)
        ids = [self.add_request(p, sp) for p, sp in zip(prompts, params)]
        out: Dict[int, List[int]] = {i: [] for i in ids}
        while self.has_work():
            for o in self.step():
                if o.request_id in out and o.token >= 0:
                    out[o.request_id].append(o.token)
        return [out[i] for i in ids]

    def generate_no_cache(self, prompts: Sequence[Sequence[int]], params=None) -> List[List[int]]:
        """The reference's generation mode without `--use_cache` (generate.py:147-190): every step recomputes the whole
        sequence (prompt + tokens so far) and keeps nothing between steps.  O(S^2) per token; kept for CLI parity and as
        an independent check of the cached path -- it samples with the same RNG streams, so it returns the same tokens."""
        from .runner import request_seed

        if params is None or isinstance(params, SamplingParams):
            params = [params or SamplingParams()] * len(prompts)
        r = self.runner
        outs: List[List[int]] = [[] for _ in prompts]
        live = list(range(len(prompts)))
        step = 0
        keep_prefix, r.prefix_enabled = r.prefix_enabled, False   # "no cache" means none: every step recomputes everything
        while live:
            seqs = []
            for i in live:
                s = SeqState(self._next_id + i, list(prompts[i]) + outs[i], params[i], eos=self.eos_token_id)
                seqs.append(s)
            # the counter RNG of request i has already produced len(outs[i]) draws
            toks = r.prefill(seqs, rng_steps=[len(outs[i]) for i in live])
            for s in seqs:
                r.release(s)
            nxt = []
            for i, t in zip(live, toks):
                outs[i].append(int(t))
                full = len(prompts[i]) + len(outs[i]) >= r.max_seq_len
                if len(outs[i]) < params[i].max_new_tokens and not full and not (self.eos_token_id is not None and t == self.eos_token_id):
                    nxt.append(i)
            live = nxt
            step += 1
        r.prefix_enabled = keep_prefix
        self._next_id += len(prompts)
        return outs
