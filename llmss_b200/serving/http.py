"""HTTP producer: `POST /generate` -> broker -> (consumer) -> broker -> response.

Parity: reference `poc-server/producer-consumer/producer_server.py:9-55` (same route, same JSON
schema, same flags in servers/producer_server.py).  Fixes of quirk Q14: every request carries a
request id and waits on its *own* reply list with a blocking pop in a worker thread, so concurrent
clients get their own answers and the event loop is never blocked by a busy-wait.
Extra routes: `GET /health`, `GET /metrics` (Prometheus text), `POST /generate_stream` (SSE),
`POST /v1/completions` (OpenAI-style request / response on the same broker path).
"""
from __future__ import annotations

import asyncio
import json
from typing import Callable, Optional

from fastapi import FastAPI, HTTPException
from fastapi.responses import PlainTextResponse, StreamingResponse
from pydantic import BaseModel

from ..utils.metrics import METRICS
from . import health as _health
from .broker import PRODUCER_QUEUE
from .protocol import GenerateRequest, GenerateResponse


class Request(BaseModel):
    prompt: str
    max_new_tokens: int = 20
    is_greedy: bool = False
    temperature: float = 1.0
    top_p: float = 0.95
    top_k: int = 50
    seed: Optional[int] = None


class Response(BaseModel):
    prompt: str
    continuation: str


def create_app(broker_factory: Callable[[], object], timeout_s: float = 600.0, down_grace_s: float = 30.0) -> FastAPI:
    """`broker_factory()` returns a client with lpush/brpop (one per worker thread)."""
    app = FastAPI(title="llmss_b200 producer")

    def _submit_and_wait(req: GenerateRequest) -> GenerateResponse:
        client = broker_factory()
        try:
            req.with_routing()
            client.lpush(PRODUCER_QUEUE, req.to_json())
            got = _health.wait_reply(client, req.reply_to, timeout_s, down_grace_s=down_grace_s)
            if got is None:
                raise TimeoutError("no response from the model server")
            return GenerateResponse.from_json(got[1])
        finally:
            close = getattr(client, "close", None)
            if close:
                close()

    @app.post("/generate")
    async def generate(request: Request) -> Response:
        METRICS.inc("http_requests_total")
        req = GenerateRequest(**request.model_dump())
        try:
            req.sampling()
        except AssertionError as e:
            raise HTTPException(status_code=422, detail=str(e))
        try:
            resp = await asyncio.get_running_loop().run_in_executor(None, _submit_and_wait, req)
        except TimeoutError as e:
            raise HTTPException(status_code=504, detail=str(e))
        except _health.ConsumerDown as e:
            raise HTTPException(status_code=503, detail=str(e))
        if resp.error:
            raise HTTPException(status_code=500, detail=resp.error)
        return Response(prompt=resp.prompt, continuation=resp.continuation)

    @app.post("/generate_stream")
    async def generate_stream(request: Request):
        req = GenerateRequest(**request.model_dump(), stream=True).with_routing()
        loop = asyncio.get_running_loop()
        client = broker_factory()
        client.lpush(PRODUCER_QUEUE, req.to_json())

        async def events():
            try:
                while True:
                    try:
                        got = await loop.run_in_executor(
                            None, lambda: _health.wait_reply(client, req.reply_to, timeout_s, down_grace_s=down_grace_s))
                    except _health.ConsumerDown as e:
                        yield f"event: error\ndata: {e}\n\n"
                        return
                    if got is None:
                        yield "event: error\ndata: timeout\n\n"
                        return
                    yield f"data: {got[1]}\n\n"
                    if json.loads(got[1]).get("done", True):
                        return
            finally:
                close = getattr(client, "close", None)
                if close:
                    close()

        return StreamingResponse(events(), media_type="text/event-stream")

    # ---- OpenAI-style completions on the same path (beyond the reference's schema; non-streaming).  `prompt` may be a
    # string or a list of strings; every prompt becomes one request to the consumer, answered as one choice.
    @app.post("/v1/completions")
    async def completions(body: dict):
        import time
        import uuid

        prompts = body.get("prompt")
        if isinstance(prompts, str):
            prompts = [prompts]
        if not isinstance(prompts, list) or not prompts or not all(isinstance(p, str) for p in prompts):
            raise HTTPException(status_code=422, detail="'prompt' must be a string or a non-empty list of strings")
        temperature = float(body.get("temperature", 1.0))
        reqs = []
        for p in prompts:
            req = GenerateRequest(prompt=p, max_new_tokens=int(body.get("max_tokens", 16)), is_greedy=temperature == 0.0,
                                  temperature=temperature if temperature > 0.0 else 1.0, top_p=float(body.get("top_p", 1.0)),
                                  top_k=int(body.get("top_k", 0)), seed=body.get("seed"))
            try:
                req.sampling()
            except AssertionError as e:
                raise HTTPException(status_code=422, detail=str(e))
            reqs.append(req)
        METRICS.inc("http_requests_total", len(reqs))
        loop = asyncio.get_running_loop()
        try:
            resps = await asyncio.gather(*[loop.run_in_executor(None, _submit_and_wait, r) for r in reqs])
        except TimeoutError as e:
            raise HTTPException(status_code=504, detail=str(e))
        except _health.ConsumerDown as e:
            raise HTTPException(status_code=503, detail=str(e))
        for r in resps:
            if r.error:
                raise HTTPException(status_code=500, detail=r.error)
        n_prompt = sum(r.prompt_tokens or 0 for r in resps)
        n_out = sum(r.num_tokens or 0 for r in resps)
        return {"id": f"cmpl-{uuid.uuid4().hex[:24]}", "object": "text_completion", "created": int(time.time()),
                "model": body.get("model", "llmss_b200"),
                "choices": [{"index": i, "text": r.continuation, "logprobs": None,
                             "finish_reason": "length" if (r.num_tokens or 0) >= reqs[i].max_new_tokens else "stop"}
                            for i, r in enumerate(resps)],
                "usage": {"prompt_tokens": n_prompt, "completion_tokens": n_out, "total_tokens": n_prompt + n_out}}

    @app.get("/health")
    async def health():
        # status: ok | consumer-down (no heartbeat) | consumer-failed | broker-unreachable; `consumer` is the heartbeat
        return _health.status(broker_factory)

    @app.get("/metrics", response_class=PlainTextResponse)
    async def metrics():
        return METRICS.render_prometheus()

    return app
