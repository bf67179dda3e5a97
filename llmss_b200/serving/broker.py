"""Request broker: a work queue speaking the Redis list protocol the reference uses.

The reference's "pubsub" is two Redis lists used as FIFO queues:
requests  `LPUSH pqueue <json>` / `LLEN pqueue` + `RPOP pqueue`   (producer_server.py:48, consumer_server.py:79-80)
responses `LPUSH squeue <json>` / `LLEN squeue` + `RPOP squeue`   (consumer_server.py:172-173, producer_server.py:51-52)

`redis-server` and the `redis` module are not installable on the target boxes, so this module ships
  * `BrokerServer`  -- a small threaded TCP server speaking the RESP2 wire protocol for the command
                       subset below, so the reference's own producer/consumer (or any redis client)
                       can talk to it unchanged;
  * `BrokerClient`  -- a socket client with redis-py's method names (`lpush`, `rpop`, `llen`, ...);
  * `InProcBroker`  -- the same interface without sockets (tests, single-process serving);
  * `connect()`     -- returns redis-py's client when that package exists (real Redis), else ours.

Beyond the reference's three commands it implements BRPOP/BLPOP (blocking pop with timeout: no
busy-wait, fixes Q14), RPUSH/LPOP, DEL, EXISTS, SET/GET, EXPIRE, PING, ECHO, FLUSHALL and answers
CLIENT/HELLO/COMMAND/SELECT handshakes of real clients.
"""
from __future__ import annotations

import socket
import socketserver
import threading
import time
from collections import defaultdict, deque
from typing import Deque, Dict, List, Optional, Tuple

PRODUCER_QUEUE = "pqueue"    # literal queue names of the reference (consumer_server.py:39-40)
SUBSCRIBER_QUEUE = "squeue"


class _Store:
    """Thread-safe lists + strings with blocking pops."""

    def __init__(self):
        self.lock = threading.Condition()
        self.lists: Dict[str, Deque[bytes]] = defaultdict(deque)
        self.strings: Dict[str, bytes] = {}
        self.expiry: Dict[str, float] = {}

    def _expire(self):
        now = time.monotonic()
        for k in [k for k, t in self.expiry.items() if t <= now]:
            self.expiry.pop(k, None)
            self.lists.pop(k, None)
            self.strings.pop(k, None)

    def push(self, key: str, values: List[bytes], left: bool) -> int:
        with self.lock:
            self._expire()
            q = self.lists[key]
            for v in values:
                q.appendleft(v) if left else q.append(v)
            self.lock.notify_all()
            return len(q)

    def pop(self, key: str, left: bool) -> Optional[bytes]:
        with self.lock:
            self._expire()
            q = self.lists.get(key)
            if not q:
                return None
            v = q.popleft() if left else q.pop()
            if not q:
                self.lists.pop(key, None)
            return v

    def bpop(self, keys: List[str], timeout: float, left: bool) -> Optional[Tuple[str, bytes]]:
        deadline = None if timeout <= 0 else time.monotonic() + timeout
        with self.lock:
            while True:
                self._expire()
                for k in keys:
                    q = self.lists.get(k)
                    if q:
                        v = q.popleft() if left else q.pop()
                        if not q:
                            self.lists.pop(k, None)
                        return k, v
                remaining = None if deadline is None else deadline - time.monotonic()
                if remaining is not None and remaining <= 0:
                    return None
                self.lock.wait(timeout=remaining if remaining is not None else 1.0)

    def llen(self, key: str) -> int:
        with self.lock:
            self._expire()
            return len(self.lists.get(key, ()))

    def delete(self, keys: List[str]) -> int:
        with self.lock:
            n = 0
            for k in keys:
                n += int(self.lists.pop(k, None) is not None) + int(self.strings.pop(k, None) is not None)
                self.expiry.pop(k, None)
            return n


# ------------------------------------------------------------------------------- in-process
class InProcBroker:
    """redis-py-like interface over an in-memory store (decode_responses=True semantics)."""

    def __init__(self, store: Optional[_Store] = None):
        self._s = store or _Store()

    @staticmethod
    def _b(v) -> bytes:
        return v if isinstance(v, bytes) else str(v).encode()

    def ping(self) -> bool:
        return True

    def lpush(self, key, *values) -> int:
        return self._s.push(key, [self._b(v) for v in values], left=True)

    def rpush(self, key, *values) -> int:
        return self._s.push(key, [self._b(v) for v in values], left=False)

    def rpop(self, key) -> Optional[str]:
        v = self._s.pop(key, left=False)
        return None if v is None else v.decode()

    def lpop(self, key) -> Optional[str]:
        v = self._s.pop(key, left=True)
        return None if v is None else v.decode()

    def brpop(self, keys, timeout: float = 0):
        keys = [keys] if isinstance(keys, str) else list(keys)
        r = self._s.bpop(keys, timeout, left=False)
        return None if r is None else (r[0], r[1].decode())

    def blpop(self, keys, timeout: float = 0):
        keys = [keys] if isinstance(keys, str) else list(keys)
        r = self._s.bpop(keys, timeout, left=True)
        return None if r is None else (r[0], r[1].decode())

    def llen(self, key) -> int:
        return self._s.llen(key)

    def delete(self, *keys) -> int:
        return self._s.delete(list(keys))

    def set(self, key, value, ex: Optional[float] = None) -> bool:
        with self._s.lock:
            self._s.strings[key] = self._b(value)
            if ex:
                self._s.expiry[key] = time.monotonic() + ex
            else:
                self._s.expiry.pop(key, None)   # SET without EX clears an earlier time to live
        return True

    def get(self, key) -> Optional[str]:
        with self._s.lock:
            self._s._expire()
            v = self._s.strings.get(key)
        return None if v is None else v.decode()

    def expire(self, key, seconds: float) -> bool:
        with self._s.lock:
            if key in self._s.lists or key in self._s.strings:
                self._s.expiry[key] = time.monotonic() + seconds
                return True
        return False

    def close(self):
        pass


# ------------------------------------------------------------------------------- RESP server
def _enc_bulk(v: Optional[bytes]) -> bytes:
    return b"$-1\r\n" if v is None else b"$%d\r\n%s\r\n" % (len(v), v)


def _enc_int(n: int) -> bytes:
    return b":%d\r\n" % n


class _Handler(socketserver.StreamRequestHandler):
    def _read_command(self) -> Optional[List[bytes]]:
        line = self.rfile.readline()
        if not line:
            return None
        line = line.rstrip(b"\r\n")
        if not line:
            return []
        if line[:1] != b"*":  # inline command
            return line.split()
        n = int(line[1:])
        args = []
        for _ in range(n):
            hdr = self.rfile.readline().rstrip(b"\r\n")
            assert hdr[:1] == b"$", "protocol error"
            ln = int(hdr[1:])
            data = self.rfile.read(ln + 2)[:-2]
            args.append(data)
        return args

    def handle(self):
        store: _Store = self.server.store  # type: ignore[attr-defined]
        while True:
            try:
                args = self._read_command()
            except Exception:
                return
            if args is None:
                return
            if not args:
                continue
            cmd = args[0].upper().decode()
            a = args[1:]
            try:
                out = self._dispatch(store, cmd, a)
            except Exception as e:  # noqa: BLE001 - never let one bad request kill the broker
                out = b"-ERR %s\r\n" % str(e).encode()[:200]
            try:
                self.wfile.write(out)
                self.wfile.flush()
            except Exception:
                return
            if cmd == "QUIT":
                return

    @staticmethod
    def _dispatch(store: _Store, cmd: str, a: List[bytes]) -> bytes:
        s = lambda b: b.decode()  # noqa: E731
        if cmd == "PING":
            return b"+PONG\r\n" if not a else _enc_bulk(a[0])
        if cmd == "ECHO":
            return _enc_bulk(a[0])
        if cmd in ("LPUSH", "RPUSH"):
            return _enc_int(store.push(s(a[0]), a[1:], left=(cmd == "LPUSH")))
        if cmd in ("RPOP", "LPOP"):
            return _enc_bulk(store.pop(s(a[0]), left=(cmd == "LPOP")))
        if cmd in ("BRPOP", "BLPOP"):
            timeout = float(a[-1])
            r = store.bpop([s(k) for k in a[:-1]], timeout, left=(cmd == "BLPOP"))
            if r is None:
                return b"*-1\r\n"
            return b"*2\r\n" + _enc_bulk(r[0].encode()) + _enc_bulk(r[1])
        if cmd == "LLEN":
            return _enc_int(store.llen(s(a[0])))
        if cmd == "DEL":
            return _enc_int(store.delete([s(k) for k in a]))
        if cmd == "EXISTS":
            with store.lock:
                return _enc_int(sum(int(s(k) in store.lists or s(k) in store.strings) for k in a))
        if cmd == "SET":
            with store.lock:
                store.strings[s(a[0])] = a[1]
                store.expiry.pop(s(a[0]), None)
                for i in range(2, len(a) - 1):
                    if a[i].upper() == b"EX":
                        store.expiry[s(a[0])] = time.monotonic() + float(a[i + 1])
            return b"+OK\r\n"
        if cmd == "GET":
            with store.lock:
                store._expire()
                return _enc_bulk(store.strings.get(s(a[0])))
        if cmd == "EXPIRE":
            with store.lock:
                k = s(a[0])
                if k in store.lists or k in store.strings:
                    store.expiry[k] = time.monotonic() + float(a[1])
                    return _enc_int(1)
            return _enc_int(0)
        if cmd in ("FLUSHALL", "FLUSHDB"):
            with store.lock:
                store.lists.clear()
                store.strings.clear()
                store.expiry.clear()
            return b"+OK\r\n"
        if cmd in ("SELECT", "QUIT", "AUTH"):
            return b"+OK\r\n"
        if cmd == "CLIENT":
            return b"+OK\r\n"
        if cmd == "COMMAND":
            return b"*0\r\n"
        if cmd == "HELLO":
            return b"-NOPROTO unsupported protocol version\r\n"  # clients fall back to RESP2
        if cmd == "INFO":
            return _enc_bulk(b"# Server\r\nredis_version:7.0.0-llmss_b200\r\n")
        return b"-ERR unknown command '%s'\r\n" % cmd.encode()


class _TCPServer(socketserver.ThreadingTCPServer):
    allow_reuse_address = True
    daemon_threads = True


class BrokerServer:
    """`BrokerServer(host, port).start()`; `port=0` picks a free port (see `.port`)."""

    def __init__(self, host: str = "127.0.0.1", port: int = 20000):
        self.store = _Store()
        self._srv = _TCPServer((host, port), _Handler)
        self._srv.store = self.store  # type: ignore[attr-defined]
        self.host, self.port = self._srv.server_address[:2]
        self._thread: Optional[threading.Thread] = None

    def start(self) -> "BrokerServer":
        self._thread = threading.Thread(target=self._srv.serve_forever, name="llmss-broker", daemon=True)
        self._thread.start()
        return self

    def serve_forever(self):
        self._srv.serve_forever()

    def stop(self):
        self._srv.shutdown()
        self._srv.server_close()


# ------------------------------------------------------------------------------- RESP client
class BrokerClient:
    """Minimal RESP2 client with redis-py's method names (``decode_responses=True`` behaviour)."""

    def __init__(self, host: str = "127.0.0.1", port: int = 20000, timeout: Optional[float] = None,
                 decode_responses: bool = True):
        self._addr = (host, int(port))
        self._decode = decode_responses
        self._lock = threading.Lock()
        self._sock = socket.create_connection(self._addr, timeout=timeout)
        self._sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        self._f = self._sock.makefile("rb")

    def _send(self, *args):
        parts = [b"*%d\r\n" % len(args)]
        for x in args:
            b = x if isinstance(x, bytes) else str(x).encode()
            parts.append(b"$%d\r\n%s\r\n" % (len(b), b))
        self._sock.sendall(b"".join(parts))

    def _read(self):
        line = self._f.readline()
        if not line:
            raise ConnectionError("broker closed the connection")
        t, rest = line[:1], line[1:].rstrip(b"\r\n")
        if t == b"+":
            return rest.decode()
        if t == b"-":
            raise RuntimeError(rest.decode())
        if t == b":":
            return int(rest)
        if t == b"$":
            n = int(rest)
            if n < 0:
                return None
            data = self._f.read(n + 2)[:-2]
            return data.decode() if self._decode else data
        if t == b"*":
            n = int(rest)
            if n < 0:
                return None
            return [self._read() for _ in range(n)]
        raise RuntimeError(f"bad RESP reply {line!r}")

    def execute(self, *args):
        with self._lock:
            self._send(*args)
            return self._read()

    def ping(self) -> bool:
        return self.execute("PING") == "PONG"

    def lpush(self, key, *values) -> int:
        return self.execute("LPUSH", key, *values)

    def rpush(self, key, *values) -> int:
        return self.execute("RPUSH", key, *values)

    def rpop(self, key):
        return self.execute("RPOP", key)

    def lpop(self, key):
        return self.execute("LPOP", key)

    def llen(self, key) -> int:
        return self.execute("LLEN", key)

    def brpop(self, keys, timeout: float = 0):
        keys = [keys] if isinstance(keys, str) else list(keys)
        r = self.execute("BRPOP", *keys, timeout)
        return None if r is None else (r[0], r[1])

    def blpop(self, keys, timeout: float = 0):
        keys = [keys] if isinstance(keys, str) else list(keys)
        r = self.execute("BLPOP", *keys, timeout)
        return None if r is None else (r[0], r[1])

    def delete(self, *keys) -> int:
        return self.execute("DEL", *keys)

    def set(self, key, value, ex: Optional[float] = None):
        return self.execute("SET", key, value, *(["EX", ex] if ex else [])) == "OK"

    def get(self, key):
        return self.execute("GET", key)

    def expire(self, key, seconds) -> bool:
        return bool(self.execute("EXPIRE", key, seconds))

    def close(self):
        try:
            self._sock.close()
        except Exception:
            pass


def connect(host: str = "127.0.0.1", port: int = 20000, prefer_redis: bool = True):
    """A list-queue client for `host:port`: redis-py if importable (real Redis), else BrokerClient."""
    if prefer_redis:
        try:
            import redis  # type: ignore

            return redis.Redis(host=host, port=int(port), decode_responses=True)
        except Exception:
            pass
    return BrokerClient(host, int(port))
