"""Tracing / profiling hooks (the reference has none: SURVEY.md section 5.1).

Two switches, both off by default so the hot path pays one attribute test per op:
  * ``LLMSS_NVTX=1``        -- NVTX ranges around every op and every layer (visible in Nsight tools);
  * ``LLMSS_PROFILE_OPS=1`` -- per-op device time with CUDA events on the launching stream
                               (eager mode only: events are recorded per launch and resolved lazily),
                               reported by ``report()`` / ``bench.py --profile-ops``.
"""
from __future__ import annotations

import os
from collections import defaultdict
from contextlib import contextmanager

import torch

NVTX = os.environ.get("LLMSS_NVTX", "0") == "1"
PROFILE = os.environ.get("LLMSS_PROFILE_OPS", "0") == "1"
ENABLED = NVTX or PROFILE
_pending = []
_totals = defaultdict(lambda: [0, 0.0])


def enable(profile: bool = True, nvtx: bool = False):
    global PROFILE, NVTX, ENABLED
    PROFILE, NVTX = profile, nvtx
    ENABLED = PROFILE or NVTX


@contextmanager
def trace(name: str):
    if not ENABLED or not torch.cuda.is_available() or torch.cuda.is_current_stream_capturing():
        yield
        return
    if NVTX:
        torch.cuda.nvtx.range_push(name)
    if PROFILE:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
    try:
        yield
    finally:
        if PROFILE:
            b.record()
            _pending.append((name, a, b))
        if NVTX:
            torch.cuda.nvtx.range_pop()


def _drain():
    if _pending:
        torch.cuda.synchronize()
        for name, a, b in _pending:
            t = _totals[name]
            t[0] += 1
            t[1] += a.elapsed_time(b) * 1e3
        _pending.clear()


def reset():
    _pending.clear()
    _totals.clear()


def report() -> dict:
    """{op: {"calls": n, "total_us": t, "avg_us": t/n}} sorted by total device time."""
    _drain()
    return {k: {"calls": n, "total_us": round(t, 1), "avg_us": round(t / max(n, 1), 2)}
            for k, (n, t) in sorted(_totals.items(), key=lambda kv: -kv[1][1])}
