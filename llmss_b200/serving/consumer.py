"""Consumer / model server: pulls requests from the broker, runs them through the Engine with
continuous batching, pushes responses back.

Parity: reference `poc-server/producer-consumer/consumer_server.py:33-173` (same queue names and
JSON bodies).  Differences, all deliberate (quirk Q14):
  * many requests in flight (continuous batching) instead of batch size 1;
  * rank 0 shares each step's *new requests* with the other TP ranks as ONE fixed-layout int32
    tensor broadcast (no pickled objects, no per-token broadcast -- sampling is replicated);
  * when idle, rank 0 blocks in BRPOP with a timeout and the other ranks block in the broadcast:
    nobody spins;
  * a malformed request produces an error response for that request only; the server lives on;
  * responses go to the request's own reply list when it names one (`reply_to`), else to `squeue`;
  * rank 0 keeps a heartbeat in the broker; when the loop dies (device watchdog trap, NCCL error, engine bug) every
    request in flight is answered with an error, the heartbeat turns to `failed` and the exception propagates so
    the launcher takes the job down (serving/health.py).
"""
from __future__ import annotations

import struct
import time
from typing import Dict, List, Optional

import torch

from ..engine.engine import Engine
from ..utils.logging import logger
from ..utils.metrics import METRICS
from . import health
from .broker import PRODUCER_QUEUE, SUBSCRIBER_QUEUE
from .protocol import GenerateRequest, GenerateResponse

_HDR = 8  # ints per request header in the plan tensor


def _f2i(x: float) -> int:
    return struct.unpack("<i", struct.pack("<f", float(x)))[0]


def _i2f(x: int) -> float:
    return struct.unpack("<f", struct.pack("<i", int(x)))[0]


class _Live:
    __slots__ = ("req", "ids", "text_len", "tokens")

    def __init__(self, req: GenerateRequest):
        self.req, self.tokens, self.text_len, self.ids = req, [], 0, 0   # ids: number of prompt tokens


class ConsumerLoop:
    def __init__(self, engine: Engine, tokenizer, broker, max_new_per_step: int = 16, plan_ints: int = 1 << 18,
                 idle_block_s: float = 0.05, poll_every: int = 4):
        self.engine, self.tok, self.broker = engine, tokenizer, broker
        self.comm = engine.comm
        self.rank = self.comm.rank
        self.max_new_per_step = max_new_per_step
        self.idle_block_s = idle_block_s
        dev = engine.core.device
        self.plan = torch.zeros(plan_ints, dtype=torch.int32, device=dev if dev.type == "cuda" else "cpu")
        self.hdr = torch.zeros(4, dtype=torch.int32, device=self.plan.device)
        # While the engine is busy the broker is polled (and the ranks synchronise on the 16-byte header) only every
        # `poll_every` steps -- the schedule is a function of the step counter, so every rank knows without asking
        # whether this step has a header; a pure decode step in between costs no collective and no host sync.
        self.poll_every = max(1, int(poll_every))
        self._step_idx = 0
        self.bytes_shared = 0     # collective payload so far (tests / serving bench)
        self.live: Dict[int, _Live] = {}
        self._next_rid = 0
        self.stop = False
        self._last_beat = 0.0
        self.reply_ttl_s = 600   # an answer nobody collects (client gave up) does not stay in the broker for ever

    # ------------------------------------------------------------------ rank 0: intake
    def _reply(self, req: GenerateRequest, resp: GenerateResponse):
        resp.request_id = req.request_id
        self.broker.lpush(req.reply_to or SUBSCRIBER_QUEUE, resp.to_json())
        if req.reply_to and resp.done is not False:
            self.broker.expire(req.reply_to, self.reply_ttl_s)

    def _beat(self, force: bool = False):
        now = time.monotonic()
        if self.rank == 0 and (force or now - self._last_beat >= health.HEARTBEAT_PERIOD_S):
            self._last_beat = now
            health.beat(self.broker, "ok", tp=self.comm.size, step=self._step_idx, in_flight=len(self.live),
                        waiting=len(self.engine.waiting))

    def _fail_all(self, exc: BaseException):
        """The loop is dying: answer what is in flight, mark the heartbeat, count it (rank 0; best effort)."""
        METRICS.inc("engine_failures")
        msg = f"{type(exc).__name__}: {exc}"[:500]
        logger.error(f"serving loop failed on rank {self.rank}: {msg}")
        if self.rank != 0:
            return
        for lv in list(self.live.values()):
            try:
                self._reply(lv.req, GenerateResponse(prompt=lv.req.prompt, continuation="", done=True,
                                                     error=f"model server failed: {msg}"))
            except Exception:  # noqa: BLE001 - the broker may be what failed
                break
        self.live.clear()
        try:
            health.beat(self.broker, "failed", ttl_s=health.FAILED_TTL_S, error=msg, step=self._step_idx)
        except Exception:  # noqa: BLE001
            pass

    def _pull(self, block: bool) -> List[tuple]:
        """[(request, token ids)], validated and tokenised (rank 0 only)."""
        out = []
        max_len = self.engine.runner.max_seq_len
        while len(out) < self.max_new_per_step:
            if block and not out:
                got = self.broker.brpop(PRODUCER_QUEUE, timeout=self.idle_block_s)
                msg = None if got is None else got[1]
            else:
                msg = self.broker.rpop(PRODUCER_QUEUE)
            if msg is None:
                break
            req = None
            try:
                req = GenerateRequest.from_json(msg)
                if req.prompt == "__shutdown__":
                    self.stop = True
                    break
                req.sampling()
                ids = self.tok(req.prompt, return_attention_mask=False, truncation=True, max_length=max_len - 1)["input_ids"]
                if len(ids) == 0:
                    raise ValueError("prompt tokenises to nothing")
                out.append((req, [int(t) for t in ids]))
            except Exception as e:  # noqa: BLE001 - isolate the failure to this request
                METRICS.inc("requests_rejected")
                logger.warning(f"rejected request: {e}")
                if req is not None:
                    self._reply(req, GenerateResponse(prompt=req.prompt, continuation="", error=str(e)))
        return out

    # ------------------------------------------------------------------ plan broadcast
    def _share(self, new: List[tuple]) -> List[tuple]:
        """Make every rank see the same [(rid, ids, SamplingParams)] list (+ stop flag).

        Wire format: a 4-int header [n_requests, stop, flat_len, 0] is broadcast first; only when n_requests > 0 a
        second broadcast carries exactly the used prefix of the plan buffer (rounded up to 256 ints), and the
        followers decode exactly ``flat_len`` ints of it."""
        items = []
        flat: List[int] = []
        if self.rank == 0:
            for req, ids in new:
                rid = self._next_rid
                self._next_rid += 1
                self.live[rid] = _Live(req)
                self.live[rid].ids = len(ids)
                sp = req.sampling()
                flat += [rid, len(ids), sp.max_new_tokens, int(sp.is_greedy), sp.top_k, _f2i(sp.temperature),
                         _f2i(sp.top_p), -1 if sp.seed is None else (sp.seed & 0x7FFFFFFF)]
                flat += ids
                items.append((rid, ids, sp))
            if len(flat) > self.plan.numel():
                raise RuntimeError("request plan exceeds the broadcast buffer")
        if self.comm.single:
            return items
        if self.rank == 0:
            self.hdr.copy_(torch.tensor([len(new), int(self.stop), len(flat), 0], dtype=torch.int32))
        self.comm.broadcast(self.hdr, src=0)
        self.bytes_shared += 16
        n, stop, flat_len, _ = self.hdr.tolist()
        self.stop = bool(stop)
        if n == 0:
            return items
        padded = min(self.plan.numel(), -(-flat_len // 256) * 256)
        if self.rank == 0:
            self.plan[:flat_len] = torch.tensor(flat, dtype=torch.int32)
        self.comm.broadcast(self.plan[:padded], src=0)
        self.bytes_shared += padded * 4
        if self.rank == 0:
            return items
        from ..engine.runner import SamplingParams

        buf = self.plan[:flat_len].tolist()
        pos = 0
        for _ in range(n):
            rid, L, mx, gr, tk, te, tp, seed = buf[pos:pos + _HDR]
            pos += _HDR
            ids = buf[pos:pos + L]
            pos += L
            items.append((rid, ids, SamplingParams(max_new_tokens=mx, is_greedy=bool(gr), top_k=tk,
                                                   temperature=_i2f(te), top_p=_i2f(tp), seed=None if seed < 0 else seed)))
        return items

    # ------------------------------------------------------------------ main loop
    def step(self) -> int:
        """One scheduling iteration; returns the number of tokens produced."""
        idle = not self.engine.has_work()
        self._step_idx += 1
        self._beat()
        if idle or self._step_idx % self.poll_every == 0:   # same decision on every rank (replicated engine state)
            new = self._pull(block=idle) if self.rank == 0 else []
            for rid, ids, sp in self._share(new):
                self.engine.add_request(ids, sp, request_id=rid)
        if not self.engine.has_work():
            return 0
        t_it = time.perf_counter()
        decode_only = not self.engine.waiting
        outs = self.engine.step()
        if self.rank != 0:
            for o in outs:
                if o.finished:
                    self.engine.stats.pop(o.request_id, None)
        if self.rank == 0:
            for o in outs:
                lv = self.live.get(o.request_id)
                if lv is None:
                    continue
                if o.token < 0:  # rejected by the engine: can never fit the KV cache
                    self._reply(lv.req, GenerateResponse(prompt=lv.req.prompt, continuation="",
                                                         error="request does not fit the KV cache (prompt + max_new_tokens too long)"))
                    del self.live[o.request_id]
                    self.engine.stats.pop(o.request_id, None)
                    continue
                lv.tokens.append(o.token)
                if lv.req.stream and not o.finished:
                    text = self.tok.decode(lv.tokens)
                    delta, lv.text_len = text[lv.text_len:], len(text)
                    if delta:
                        self._reply(lv.req, GenerateResponse(prompt=lv.req.prompt, continuation="", delta=delta, done=False))
                if o.finished:
                    st = self.engine.stats.get(o.request_id)
                    text = self.tok.decode(lv.tokens)
                    self._reply(lv.req, GenerateResponse(
                        prompt=lv.req.prompt, continuation=text, done=True, num_tokens=len(lv.tokens), prompt_tokens=lv.ids,
                        ttft_ms=None if st is None or st.ttft is None else round(1e3 * st.ttft, 3),
                        tpot_ms=None if st is None or st.tpot is None else round(1e3 * st.tpot, 3)))
                    del self.live[o.request_id]
                    self.engine.stats.pop(o.request_id, None)
        if decode_only:  # a whole serving iteration around a decode step (poll + share + step + replies)
            METRICS.observe("consumer_step_seconds", time.perf_counter() - t_it)
        return len(outs)

    def run_forever(self):
        if self.rank == 0:
            logger.info(f"{type(self.engine.model).__name__} setup is done.")
        try:
            self._beat(force=True)
            while not self.stop:
                self.step()
            # drain what is in flight before leaving
            while self.engine.has_work():
                self.step()
        except BaseException as e:
            self._fail_all(e)
            raise
        if self.rank == 0:
            try:
                health.clear(self.broker)
            except Exception:  # noqa: BLE001 - the broker may already be gone at shutdown
                pass
