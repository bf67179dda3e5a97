"""Wire schema shared by the HTTP producer, the gRPC service and the consumer.

Parity: request / response bodies of the reference's `POST /generate`
(producer_server.py:9-20): ``{prompt, max_new_tokens, is_greedy, temperature, top_p, top_k}``
-> ``{prompt, continuation}``.  Additions (all optional, ignored by reference-style peers):
``request_id`` + ``reply_to`` give every request its own reply list, so concurrent clients can no
longer steal each other's answers (quirk Q14); ``stream`` asks for incremental deltas; ``seed``.
"""
from __future__ import annotations

import json
import uuid
from dataclasses import asdict, dataclass, field
from typing import Optional

from ..engine.runner import SamplingParams


@dataclass
class GenerateRequest:
    prompt: str
    max_new_tokens: int = 20
    is_greedy: bool = False
    temperature: float = 1.0
    top_p: float = 0.95
    top_k: int = 50
    seed: Optional[int] = None
    stream: bool = False
    request_id: Optional[str] = None
    reply_to: Optional[str] = None

    def sampling(self) -> SamplingParams:
        return SamplingParams(max_new_tokens=int(self.max_new_tokens), is_greedy=bool(self.is_greedy),
                              temperature=float(self.temperature), top_p=float(self.top_p), top_k=int(self.top_k),
                              seed=None if self.seed is None else int(self.seed)).validate()

    def with_routing(self) -> "GenerateRequest":
        if self.request_id is None:
            self.request_id = uuid.uuid4().hex
        if self.reply_to is None:
            self.reply_to = f"squeue:{self.request_id}"
        return self

    def to_json(self) -> str:
        return json.dumps({k: v for k, v in asdict(self).items() if v is not None}, ensure_ascii=False)

    @classmethod
    def from_json(cls, text: str) -> "GenerateRequest":
        d = json.loads(text)
        known = {k: d[k] for k in cls.__dataclass_fields__ if k in d}
        if "prompt" not in known:
            raise ValueError("request has no 'prompt'")
        return cls(**known)


@dataclass
class GenerateResponse:
    prompt: str
    continuation: str
    request_id: Optional[str] = None
    error: Optional[str] = None
    done: bool = True
    delta: Optional[str] = None
    num_tokens: Optional[int] = None
    prompt_tokens: Optional[int] = None
    ttft_ms: Optional[float] = None
    tpot_ms: Optional[float] = None

    def to_json(self) -> str:
        return json.dumps({k: v for k, v in asdict(self).items() if v is not None}, ensure_ascii=False)

    @classmethod
    def from_json(cls, text: str) -> "GenerateResponse":
        d = json.loads(text)
        return cls(**{k: d[k] for k in cls.__dataclass_fields__ if k in d})
