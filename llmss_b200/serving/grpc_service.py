"""gRPC front end.

The reference's README advertises gRPC but ships no gRPC code, proto or import (SURVEY.md
section 0); this defines it.  `grpc_tools` / `protoc` are not available on the target boxes, so the
messages are built at run time from a `FileDescriptorProto` (equivalent .proto below) and the
service is registered with `grpc.method_handlers_generic_handler`.

    syntax = "proto3";  package llmss;
    message GenerateRequest  { string prompt = 1; int32 max_new_tokens = 2; bool is_greedy = 3;
                               float temperature = 4; float top_p = 5; int32 top_k = 6; int32 seed = 7; }
    message GenerateResponse { string prompt = 1; string continuation = 2; string delta = 3; bool done = 4;
                               string error = 5; int32 num_tokens = 6; float ttft_ms = 7; float tpot_ms = 8; }
    message HealthRequest {}   message HealthResponse { string status = 1; }
    service LLMSS { rpc Generate(GenerateRequest) returns (GenerateResponse);
                    rpc GenerateStream(GenerateRequest) returns (stream GenerateResponse);
                    rpc Health(HealthRequest) returns (HealthResponse); }

Like the HTTP producer, the service is a *front end*: it forwards requests through the broker to the
tensor-parallel consumer and mirrors the HTTP schema.
"""
from __future__ import annotations

from concurrent import futures
from typing import Callable

import grpc
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

from ..utils.metrics import METRICS
from . import health as _health
from . import protocol
from .broker import PRODUCER_QUEUE

_T = descriptor_pb2.FieldDescriptorProto
SERVICE = "llmss.LLMSS"


def _build_messages():
    fd = descriptor_pb2.FileDescriptorProto(name="llmss_b200/llmss.proto", package="llmss", syntax="proto3")

    def msg(name, fields):
        m = fd.message_type.add(name=name)
        for i, (fname, ftype) in enumerate(fields, start=1):
            m.field.add(name=fname, number=i, type=ftype, label=_T.LABEL_OPTIONAL)

    msg("GenerateRequest", [("prompt", _T.TYPE_STRING), ("max_new_tokens", _T.TYPE_INT32), ("is_greedy", _T.TYPE_BOOL),
                            ("temperature", _T.TYPE_FLOAT), ("top_p", _T.TYPE_FLOAT), ("top_k", _T.TYPE_INT32),
                            ("seed", _T.TYPE_INT32)])
    msg("GenerateResponse", [("prompt", _T.TYPE_STRING), ("continuation", _T.TYPE_STRING), ("delta", _T.TYPE_STRING),
                             ("done", _T.TYPE_BOOL), ("error", _T.TYPE_STRING), ("num_tokens", _T.TYPE_INT32),
                             ("ttft_ms", _T.TYPE_FLOAT), ("tpot_ms", _T.TYPE_FLOAT)])
    msg("HealthRequest", [])
    msg("HealthResponse", [("status", _T.TYPE_STRING)])
    svc = fd.service.add(name="LLMSS")
    svc.method.add(name="Generate", input_type=".llmss.GenerateRequest", output_type=".llmss.GenerateResponse")
    svc.method.add(name="GenerateStream", input_type=".llmss.GenerateRequest", output_type=".llmss.GenerateResponse",
                   server_streaming=True)
    svc.method.add(name="Health", input_type=".llmss.HealthRequest", output_type=".llmss.HealthResponse")
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    get = lambda n: message_factory.GetMessageClass(pool.FindMessageTypeByName(f"llmss.{n}"))  # noqa: E731
    return {n: get(n) for n in ("GenerateRequest", "GenerateResponse", "HealthRequest", "HealthResponse")}, fd


MESSAGES, FILE_DESCRIPTOR = _build_messages()
GenerateRequestPB, GenerateResponsePB = MESSAGES["GenerateRequest"], MESSAGES["GenerateResponse"]
HealthRequestPB, HealthResponsePB = MESSAGES["HealthRequest"], MESSAGES["HealthResponse"]


def _to_internal(pb, stream: bool) -> protocol.GenerateRequest:
    # proto3 scalars default to 0: treat unset sampling fields as the reference CLI defaults
    return protocol.GenerateRequest(
        prompt=pb.prompt, max_new_tokens=pb.max_new_tokens or 20, is_greedy=pb.is_greedy,
        temperature=pb.temperature or 1.0, top_p=pb.top_p or 0.95, top_k=pb.top_k, seed=(pb.seed or None), stream=stream)


def _to_pb(r: protocol.GenerateResponse):
    return GenerateResponsePB(prompt=r.prompt, continuation=r.continuation or "", delta=r.delta or "", done=bool(r.done),
                              error=r.error or "", num_tokens=r.num_tokens or 0, ttft_ms=r.ttft_ms or 0.0,
                              tpot_ms=r.tpot_ms or 0.0)


class LLMSSServicer:
    def __init__(self, broker_factory: Callable[[], object], timeout_s: float = 600.0, down_grace_s: float = 30.0):
        self.broker_factory, self.timeout_s, self.down_grace_s = broker_factory, timeout_s, down_grace_s

    def _submit(self, pb, stream: bool, context):
        req = _to_internal(pb, stream)
        try:
            req.sampling()
        except AssertionError as e:
            context.abort(grpc.StatusCode.INVALID_ARGUMENT, str(e))
        req.with_routing()
        client = self.broker_factory()
        client.lpush(PRODUCER_QUEUE, req.to_json())
        return req, client

    def Generate(self, pb, context):
        METRICS.inc("grpc_requests_total")
        req, client = self._submit(pb, False, context)
        try:
            got = _health.wait_reply(client, req.reply_to, self.timeout_s, down_grace_s=self.down_grace_s)
        except _health.ConsumerDown as e:
            context.abort(grpc.StatusCode.UNAVAILABLE, str(e))
        finally:
            getattr(client, "close", lambda: None)()
        if got is None:
            context.abort(grpc.StatusCode.DEADLINE_EXCEEDED, "no response from the model server")
        resp = protocol.GenerateResponse.from_json(got[1])
        if resp.error:
            context.abort(grpc.StatusCode.INTERNAL, resp.error)
        return _to_pb(resp)

    def GenerateStream(self, pb, context):
        METRICS.inc("grpc_requests_total")
        req, client = self._submit(pb, True, context)
        try:
            while True:
                try:
                    got = _health.wait_reply(client, req.reply_to, self.timeout_s, down_grace_s=self.down_grace_s)
                except _health.ConsumerDown as e:
                    context.abort(grpc.StatusCode.UNAVAILABLE, str(e))
                if got is None:
                    context.abort(grpc.StatusCode.DEADLINE_EXCEEDED, "no response from the model server")
                resp = protocol.GenerateResponse.from_json(got[1])
                yield _to_pb(resp)
                if resp.done:
                    return
        finally:
            getattr(client, "close", lambda: None)()

    def Health(self, pb, context):
        return HealthResponsePB(status=_health.status(self.broker_factory)["status"])


def add_servicer(server: grpc.Server, servicer: LLMSSServicer):
    ser = lambda m: m.SerializeToString()  # noqa: E731
    handlers = {
        "Generate": grpc.unary_unary_rpc_method_handler(servicer.Generate, GenerateRequestPB.FromString, ser),
        "GenerateStream": grpc.unary_stream_rpc_method_handler(servicer.GenerateStream, GenerateRequestPB.FromString, ser),
        "Health": grpc.unary_unary_rpc_method_handler(servicer.Health, HealthRequestPB.FromString, ser),
    }
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(SERVICE, handlers),))


def serve(broker_factory, host: str = "127.0.0.1", port: int = 50051, max_workers: int = 32, timeout_s: float = 600.0,
          down_grace_s: float = 30.0):
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers))
    add_servicer(server, LLMSSServicer(broker_factory, timeout_s, down_grace_s))
    bound = server.add_insecure_port(f"{host}:{port}")
    server.start()
    return server, bound


class Client:
    """Tiny client for the service above (tests, examples)."""

    def __init__(self, target: str):
        self.channel = grpc.insecure_channel(target)
        ser = lambda m: m.SerializeToString()  # noqa: E731
        self._gen = self.channel.unary_unary(f"/{SERVICE}/Generate", ser, GenerateResponsePB.FromString)
        self._stream = self.channel.unary_stream(f"/{SERVICE}/GenerateStream", ser, GenerateResponsePB.FromString)
        self._health = self.channel.unary_unary(f"/{SERVICE}/Health", ser, HealthResponsePB.FromString)

    def generate(self, prompt: str, timeout: float = 600.0, **kw):
        return self._gen(GenerateRequestPB(prompt=prompt, **kw), timeout=timeout)

    def generate_stream(self, prompt: str, timeout: float = 600.0, **kw):
        return self._stream(GenerateRequestPB(prompt=prompt, **kw), timeout=timeout)

    def health(self, timeout: float = 10.0) -> str:
        return self._health(HealthRequestPB(), timeout=timeout).status

    def close(self):
        self.channel.close()
