"""gRPC front end: `python servers/grpc_server.py --port 50051 --redis_host 127.0.0.1 --redis_port 20000`."""
import os
import sys
from argparse import ArgumentParser

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from llmss_b200.serving.broker import connect
    from llmss_b200.serving.grpc_service import serve

    p = ArgumentParser()
    p.add_argument("--host", type=str, default="127.0.0.1")
    p.add_argument("--port", type=int, default=50051)
    p.add_argument("--redis_host", type=str, default="127.0.0.1")
    p.add_argument("--redis_port", type=int, default=20000)
    a = p.parse_args()
    server, port = serve(lambda: connect(a.redis_host, a.redis_port), a.host, a.port)
    print(f"llmss_b200 gRPC service llmss.LLMSS on {a.host}:{port}", flush=True)
    server.wait_for_termination()


if __name__ == "__main__":
    main()
