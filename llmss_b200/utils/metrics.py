"""Serving metrics: counters, gauges and latency reservoirs (TTFT / TPOT / step time).

The reference has no observability beyond one wall-clock print that includes model load
(generate.py:193-194).  This registry is process-local, lock-protected,
exports Prometheus text (`render_prometheus`) for the HTTP `/metrics` route, and answers
quantile queries (`quantile("ttft_seconds", 0.5)`) for BASELINE's p50 TTFT / TPOT.
"""
from __future__ import annotations

import threading
from collections import defaultdict, deque
from typing import Deque, Dict


class Metrics:
    def __init__(self, reservoir: int = 65536):
        self._lock = threading.Lock()
        self._counters: Dict[str, float] = defaultdict(float)
        self._gauges: Dict[str, float] = {}
        self._obs: Dict[str, Deque[float]] = defaultdict(lambda: deque(maxlen=reservoir))
        self._obs_sum: Dict[str, float] = defaultdict(float)
        self._obs_n: Dict[str, int] = defaultdict(int)

    def inc(self, name: str, value: float = 1.0):
        with self._lock:
            self._counters[name] += value

    def set(self, name: str, value: float):
        with self._lock:
            self._gauges[name] = value

    def observe(self, name: str, value):
        if value is None:
            return
        with self._lock:
            self._obs[name].append(float(value))
            self._obs_sum[name] += float(value)
            self._obs_n[name] += 1

    def counter(self, name: str) -> float:
        return self._counters.get(name, 0.0)

    def quantile(self, name: str, q: float):
        with self._lock:
            xs = sorted(self._obs.get(name, ()))
        if not xs:
            return None
        return xs[min(len(xs) - 1, int(q * len(xs)))]

    def snapshot(self) -> dict:
        out = {"counters": dict(self._counters), "gauges": dict(self._gauges), "latency": {}}
        for k in list(self._obs):
            out["latency"][k] = {"n": self._obs_n[k], "mean": self._obs_sum[k] / max(1, self._obs_n[k]),
                                 "p50": self.quantile(k, 0.5), "p90": self.quantile(k, 0.9), "p99": self.quantile(k, 0.99)}
        return out

    def reset(self):
        with self._lock:
            self._counters.clear()
            self._gauges.clear()
            self._obs.clear()
            self._obs_sum.clear()
            self._obs_n.clear()

    def render_prometheus(self, prefix: str = "llmss_b200_") -> str:
        lines = []
        for k, v in sorted(self._counters.items()):
            lines += [f"# TYPE {prefix}{k} counter", f"{prefix}{k} {v}"]
        for k, v in sorted(self._gauges.items()):
            lines += [f"# TYPE {prefix}{k} gauge", f"{prefix}{k} {v}"]
        for k in sorted(self._obs):
            lines.append(f"# TYPE {prefix}{k} summary")
            for q in (0.5, 0.9, 0.99):
                v = self.quantile(k, q)
                if v is not None:
                    lines.append(f'{prefix}{k}{{quantile="{q}"}} {v}')
            lines += [f"{prefix}{k}_sum {self._obs_sum[k]}", f"{prefix}{k}_count {self._obs_n[k]}"]
        return "\n".join(lines) + "\n"


METRICS = Metrics()
