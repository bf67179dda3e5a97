"""Stand-alone broker: `python servers/broker_server.py --port 20000` (replaces `redis-server --port 20000`
of the reference recipe, poc-server/producer-consumer/README.md:9-17)."""
import os
import sys
from argparse import ArgumentParser

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llmss_b200.serving.broker import BrokerServer  # noqa: E402


def main():
    p = ArgumentParser()
    p.add_argument("--host", type=str, default="127.0.0.1")
    p.add_argument("--port", type=int, default=20000)
    a = p.parse_args()
    srv = BrokerServer(a.host, a.port)
    print(f"llmss_b200 broker listening on {srv.host}:{srv.port} (RESP: LPUSH/RPOP/LLEN/BRPOP ...)", flush=True)
    srv.serve_forever()


if __name__ == "__main__":
    main()
