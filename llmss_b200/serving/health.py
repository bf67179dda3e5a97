"""Failure detection between the front ends and the tensor-parallel consumer.

The reference has none: its producer polls `squeue` forever (`producer_server.py:45-55`) and a dead consumer
leaves every client hanging.  Here rank 0 of the consumer keeps a heartbeat key in the broker (SET with expiry,
renewed about once a second, idle or busy); the front ends' `/health` and `Health` report it, and a client waiting
for its answer gives up as soon as the heartbeat says the engine failed or has been missing for `down_grace_s`,
instead of waiting for the full request timeout.

What ends up in the heartbeat's `state`:
  ok      -- written by the serving loop (with step counter, requests in flight, TP width);
  failed  -- written once when the loop dies on an exception: a device-side watchdog trap (`SpinGuard` in
             csrc/common.cuh, a peer that never arrived), an NCCL error or an engine bug.  Every request in
             flight has been answered with an error before this is written; the process then exits non-zero so
             the launcher (torchrun) tears down the other ranks and a supervisor can restart the job;
  (absent)-- never started, shut down cleanly, or died without a chance to say so (SIGKILL): the key expires.
"""
from __future__ import annotations

import json
import time
from typing import Optional

HEARTBEAT_KEY = "llmss:consumer:heartbeat"
HEARTBEAT_TTL_S = 10      # whole seconds: a real Redis takes an integer EX
HEARTBEAT_PERIOD_S = 1.0
FAILED_TTL_S = 120


class ConsumerDown(RuntimeError):
    """The model server is not there to answer (failed, or no heartbeat within the grace period)."""


def beat(broker, state: str = "ok", ttl_s: float = HEARTBEAT_TTL_S, **info) -> None:
    broker.set(HEARTBEAT_KEY, json.dumps({"state": state, "ts": time.time(), **info}), ex=max(1, int(ttl_s)))


def clear(broker) -> None:
    broker.delete(HEARTBEAT_KEY)


def read(broker) -> Optional[dict]:
    raw = broker.get(HEARTBEAT_KEY)
    if raw is None:
        return None
    try:
        return json.loads(raw)
    except (TypeError, ValueError):
        return None


def status(broker_factory) -> dict:
    """{'status': 'ok' | 'consumer-down' | 'consumer-failed' | 'broker-unreachable', 'consumer': heartbeat or None}"""
    client = None
    try:
        client = broker_factory()
        if not client.ping():
            return {"status": "broker-unreachable", "consumer": None}
        hb = read(client)
    except Exception:  # noqa: BLE001 - any transport error means the broker is not usable
        return {"status": "broker-unreachable", "consumer": None}
    finally:
        close = getattr(client, "close", None)
        if close:
            close()
    if hb is None:
        return {"status": "consumer-down", "consumer": None}
    hb["age_s"] = round(max(0.0, time.time() - float(hb.get("ts", 0.0))), 3)
    return {"status": "ok" if hb.get("state") == "ok" else "consumer-failed", "consumer": hb}


def wait_reply(client, key: str, timeout_s: float, poll_s: float = 2.0, down_grace_s: float = 30.0):
    """BRPOP `key` for up to `timeout_s`, in slices of `poll_s`; between slices look at the heartbeat.

    Returns the (key, value) pair, or None after `timeout_s`.  Raises ConsumerDown when the heartbeat reports a
    failure, or when none has been seen for `down_grace_s` (a consumer that is still loading its weights has none
    either, hence the grace period instead of an immediate error)."""
    t0 = time.monotonic()
    last_seen = t0
    while True:
        remaining = timeout_s - (time.monotonic() - t0)
        if remaining <= 0:
            return None
        got = client.brpop(key, timeout=max(0.05, min(poll_s, remaining)))
        if got is not None:
            return got
        hb = read(client)
        now = time.monotonic()
        if hb is not None and hb.get("state") == "ok":
            last_seen = now
        elif hb is not None:
            # the failing consumer answers its in-flight requests before it writes the marker: look once more
            got = client.rpop(key)
            if got is not None:
                return key, got
            raise ConsumerDown(f"model server failed: {hb.get('error', 'unknown error')}")
        elif now - last_seen > down_grace_s:
            raise ConsumerDown(f"no heartbeat from the model server for {down_grace_s:.0f} s")
