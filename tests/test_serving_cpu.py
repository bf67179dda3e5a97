"""Serving stack on CPU: broker wire protocol, consumer loop with continuous batching, HTTP and gRPC."""
import json
import threading
import time

import pytest
import torch

from llmss_b200.engine.engine import Engine
from llmss_b200.serving.broker import (PRODUCER_QUEUE, SUBSCRIBER_QUEUE, BrokerClient, BrokerServer, InProcBroker)
from llmss_b200.serving.consumer import ConsumerLoop
from llmss_b200.serving.protocol import GenerateRequest, GenerateResponse
from llmss_b200.utils.checkpoint import write_random_checkpoint


class ByteTokenizer:
    """Stand-in tokenizer (no tokenizer files offline): one token per byte, shifted past specials."""

    def __call__(self, text, return_attention_mask=False, truncation=True, max_length=None, **_):
        ids = [3 + b for b in text.encode()]
        if max_length:
            ids = ids[-max_length:]
        return {"input_ids": ids}

    def decode(self, ids):
        return bytes((max(0, i - 3)) % 256 for i in ids).decode("latin-1")


@pytest.fixture(scope="module")
def engine(tmp_path_factory):
    path = write_random_checkpoint(tmp_path_factory.mktemp("ckpt"), "gptj-tiny", dtype=torch.float32, device="cpu", std=0.05)
    return Engine.from_pretrained(str(path), device="cpu", max_batch=4, max_seq_len=96)


def test_resp_wire_protocol_matches_reference_usage():
    srv = BrokerServer(port=0).start()
    try:
        prod, cons = BrokerClient(srv.host, srv.port), BrokerClient(srv.host, srv.port)
        assert prod.ping()
        # reference protocol: LPUSH pqueue / LLEN + RPOP (FIFO through the opposite end)
        prod.lpush(PRODUCER_QUEUE, json.dumps({"prompt": "a"}))
        prod.lpush(PRODUCER_QUEUE, json.dumps({"prompt": "b"}))
        assert cons.llen(PRODUCER_QUEUE) == 2
        assert json.loads(cons.rpop(PRODUCER_QUEUE))["prompt"] == "a"
        assert json.loads(cons.rpop(PRODUCER_QUEUE))["prompt"] == "b"
        assert cons.rpop(PRODUCER_QUEUE) is None and cons.llen(PRODUCER_QUEUE) == 0
        # blocking pop wakes up when another connection pushes; times out otherwise
        t0 = time.time()
        assert cons.brpop("nothing", timeout=0.2) is None and time.time() - t0 >= 0.15
        threading.Timer(0.1, lambda: prod.lpush(SUBSCRIBER_QUEUE, "x")).start()
        assert cons.brpop(SUBSCRIBER_QUEUE, timeout=5) == (SUBSCRIBER_QUEUE, "x")
        assert prod.set("k", "v") and cons.get("k") == "v" and cons.delete("k") == 1 and cons.get("k") is None
        with pytest.raises(RuntimeError):
            prod.execute("NOSUCHCOMMAND")
        assert prod.ping()  # an error does not poison the connection
        # unicode payloads survive (the reference README posts Korean prompts)
        prod.lpush("u", "안녕하세요")
        assert cons.rpop("u") == "안녕하세요"
    finally:
        srv.stop()


def _serve(engine, broker):
    loop = ConsumerLoop(engine, ByteTokenizer(), broker, idle_block_s=0.02)
    t = threading.Thread(target=loop.run_forever, daemon=True)
    t.start()
    return loop, t


def test_consumer_batches_requests_and_isolates_errors(engine):
    broker = InProcBroker()
    loop, t = _serve(engine, broker)
    try:
        reqs = [GenerateRequest(prompt=f"hello {i}", max_new_tokens=4 + i, is_greedy=True).with_routing() for i in range(6)]
        for r in reqs:
            broker.lpush(PRODUCER_QUEUE, r.to_json())
        broker.lpush(PRODUCER_QUEUE, "{not json")                                     # malformed: dropped
        bad = GenerateRequest(prompt="x", temperature=5.0).with_routing()             # invalid: error reply
        broker.lpush(PRODUCER_QUEUE, bad.to_json())
        legacy = json.dumps({"prompt": "legacy", "max_new_tokens": 3, "is_greedy": True, "temperature": 1.0,
                             "top_p": 0.95, "top_k": 50})                             # reference-style, no id
        broker.lpush(PRODUCER_QUEUE, legacy)
        for i, r in enumerate(reqs):
            got = broker.brpop(r.reply_to, timeout=60)
            assert got is not None
            resp = GenerateResponse.from_json(got[1])
            assert resp.request_id == r.request_id and resp.prompt == r.prompt and resp.error is None
            assert resp.num_tokens == 4 + i and resp.done
        err = GenerateResponse.from_json(broker.brpop(bad.reply_to, timeout=60)[1])
        assert err.error and "temperature" in err.error
        leg = json.loads(broker.brpop(SUBSCRIBER_QUEUE, timeout=60)[1])
        assert leg["prompt"] == "legacy" and "continuation" in leg
        # same prompt greedy twice -> same continuation (determinism through the batcher)
        a, b = (GenerateRequest(prompt="same", max_new_tokens=6, is_greedy=True).with_routing() for _ in range(2))
        broker.lpush(PRODUCER_QUEUE, a.to_json())
        broker.lpush(PRODUCER_QUEUE, b.to_json())
        ra = GenerateResponse.from_json(broker.brpop(a.reply_to, timeout=60)[1])
        rb = GenerateResponse.from_json(broker.brpop(b.reply_to, timeout=60)[1])
        assert ra.continuation == rb.continuation
    finally:
        loop.stop = True
        t.join(timeout=10)


def test_http_and_grpc_front_ends(engine):
    from fastapi.testclient import TestClient

    from llmss_b200.serving import grpc_service
    from llmss_b200.serving.http import create_app

    srv = BrokerServer(port=0).start()
    loop, t = _serve(engine, BrokerClient(srv.host, srv.port))
    factory = lambda: BrokerClient(srv.host, srv.port)  # noqa: E731
    try:
        client = TestClient(create_app(factory, timeout_s=60))
        body = {"prompt": "hi there", "max_new_tokens": 5, "is_greedy": True, "temperature": 1.0, "top_p": 0.95, "top_k": 50}
        r = client.post("/generate", json=body)
        assert r.status_code == 200 and set(r.json()) == {"prompt", "continuation"} and r.json()["prompt"] == "hi there"
        assert client.post("/generate", json={**body, "top_p": 1.5}).status_code == 422
        assert client.get("/health").json()["status"] == "ok"
        assert "llmss_b200_requests_total" in client.get("/metrics").text
        # OpenAI-style completions on the same path: greedy at temperature 0, one choice per prompt, token usage
        oa = client.post("/v1/completions", json={"model": "tiny", "prompt": ["hi there", "p1"], "max_tokens": 5, "temperature": 0})
        assert oa.status_code == 200, oa.text
        oa = oa.json()
        assert oa["object"] == "text_completion" and [c["index"] for c in oa["choices"]] == [0, 1]
        assert oa["choices"][0]["text"] == r.json()["continuation"] and oa["choices"][0]["finish_reason"] == "length"
        assert oa["usage"] == {"prompt_tokens": len("hi there") + len("p1"), "completion_tokens": 10, "total_tokens": 20}
        assert client.post("/v1/completions", json={"prompt": 3}).status_code == 422
        # concurrent clients each get their own answer (reference quirk Q14)
        outs = {}

        def call(i):
            outs[i] = client.post("/generate", json={**body, "prompt": f"p{i}", "max_new_tokens": 3 + i}).json()

        th = [threading.Thread(target=call, args=(i,)) for i in range(5)]
        [x.start() for x in th]
        [x.join() for x in th]
        assert all(outs[i]["prompt"] == f"p{i}" for i in range(5))

        server, port = grpc_service.serve(factory, port=0)
        try:
            gc = grpc_service.Client(f"127.0.0.1:{port}")
            assert gc.health() == "ok"
            resp = gc.generate("hi there", max_new_tokens=5, is_greedy=True)
            assert resp.prompt == "hi there" and resp.continuation == r.json()["continuation"] and resp.num_tokens == 5
            chunks = list(gc.generate_stream("hi there", max_new_tokens=5, is_greedy=True))
            assert chunks[-1].done and chunks[-1].continuation == resp.continuation
            import grpc

            with pytest.raises(grpc.RpcError) as ei:
                gc.generate("x", temperature=3.0)
            assert ei.value.code() == grpc.StatusCode.INVALID_ARGUMENT
            gc.close()
        finally:
            server.stop(0)
    finally:
        loop.stop = True
        t.join(timeout=10)
        srv.stop()


def test_failure_detection_heartbeat_and_error_replies(engine):
    """a dying serving loop answers what is in flight, marks its heartbeat `failed` and re-raises; the
    front ends report the consumer's state and stop waiting for a consumer that is not there."""
    from fastapi.testclient import TestClient

    from llmss_b200.serving import health
    from llmss_b200.serving.http import create_app

    broker = InProcBroker()
    factory = lambda: broker  # noqa: E731
    app = TestClient(create_app(factory, timeout_s=20, down_grace_s=0.3))
    assert health.status(factory)["status"] == "consumer-down"          # nobody has ever said hello
    assert app.get("/health").json()["status"] == "consumer-down"
    t0 = time.monotonic()
    r = app.post("/generate", json={"prompt": "anyone?", "max_new_tokens": 3, "is_greedy": True})
    assert r.status_code == 503 and "heartbeat" in r.json()["detail"] and time.monotonic() - t0 < 15
    assert broker.rpop(PRODUCER_QUEUE) is not None                       # drop the unanswered request again

    loop = ConsumerLoop(engine, ByteTokenizer(), broker, idle_block_s=0.02)
    raised = []

    def run():
        try:
            loop.run_forever()
        except RuntimeError as e:
            raised.append(e)

    real_step, calls = engine.step, [0]

    def dying_step():
        calls[0] += 1
        if calls[0] == 3:      # prefill and one decode step run, then the "device" fails
            raise RuntimeError("CUDA error: unspecified launch failure (device watchdog: peer flag wait timed out)")
        return real_step()

    engine.step = dying_step
    t = threading.Thread(target=run, daemon=True)
    try:
        t.start()
        for _ in range(200):
            if health.status(factory)["status"] == "ok":
                break
            time.sleep(0.01)
        hb = health.status(factory)
        assert hb["status"] == "ok" and hb["consumer"]["tp"] == 1 and hb["consumer"]["age_s"] < 5
        reqs = [GenerateRequest(prompt=f"doomed {i}", max_new_tokens=30, is_greedy=True).with_routing() for i in range(3)]
        for q in reqs:
            broker.lpush(PRODUCER_QUEUE, q.to_json())
        for q in reqs:
            got = health.wait_reply(broker, q.reply_to, timeout_s=30, poll_s=0.05)
            resp = GenerateResponse.from_json(got[1])
            assert resp.request_id == q.request_id and resp.error and "device watchdog" in resp.error
        t.join(timeout=10)
        assert not t.is_alive() and raised and "watchdog" in str(raised[0])
        st = health.status(factory)
        assert st["status"] == "consumer-failed" and "watchdog" in st["consumer"]["error"]
        # a request that arrives after the failure is refused at once instead of waiting out its timeout
        t0 = time.monotonic()
        r = app.post("/generate", json={"prompt": "too late", "max_new_tokens": 3, "is_greedy": True})
        assert r.status_code == 503 and "model server failed" in r.json()["detail"] and time.monotonic() - t0 < 10
    finally:
        engine.step = real_step
        loop.stop = True
        dropped = engine.abort_all()          # leave the shared engine clean: blocks back, nothing bound
        assert len(dropped) == 3 and not engine.has_work()
        assert engine.runner._outstanding == 0


def test_wait_reply_timeout_grace_and_late_answer():
    """serving/health.wait_reply: plain timeout while the consumer is alive, ConsumerDown after the grace period without
    a heartbeat, and an answer that was pushed just before the `failed` marker is still delivered."""
    from llmss_b200.serving import health

    broker = InProcBroker()
    health.beat(broker, "ok", tp=1)
    t0 = time.monotonic()
    assert health.wait_reply(broker, "nobody", timeout_s=0.3, poll_s=0.05) is None          # alive, just slow
    assert 0.25 <= time.monotonic() - t0 < 3
    health.clear(broker)
    with pytest.raises(health.ConsumerDown, match="no heartbeat"):
        health.wait_reply(broker, "nobody", timeout_s=5, poll_s=0.05, down_grace_s=0.2)
    broker.lpush("mine", "answer")
    health.beat(broker, "failed", error="boom")
    assert health.wait_reply(broker, "mine", timeout_s=5, poll_s=0.05) == ("mine", "answer")
    with pytest.raises(health.ConsumerDown, match="boom"):
        health.wait_reply(broker, "mine", timeout_s=5, poll_s=0.05)
    # SET without EX clears an earlier time to live (Redis semantics)
    broker.set("k", "v", ex=1)
    broker.set("k", "w")
    time.sleep(1.2)
    assert broker.get("k") == "w"
