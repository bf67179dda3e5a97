"""HTTP producer -- same flags as the reference (producer_server.py:23-32).

    python servers/producer_server.py --fastapi_host 127.0.0.1 --fastapi_port 8000 --redis_host 127.0.0.1 --redis_port 20000
"""
import os
import sys
from argparse import ArgumentParser

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def get_args():
    parser = ArgumentParser()
    producer_group = parser.add_argument_group("producer")
    producer_group.add_argument("--fastapi_host", type=str, default="127.0.0.1")
    producer_group.add_argument("--fastapi_port", type=int, default=8000)
    broker_group = parser.add_argument_group("broker")
    broker_group.add_argument("--redis_host", type=str, default="127.0.0.1")
    broker_group.add_argument("--redis_port", type=int, default=20000)
    return parser.parse_args()


def main():
    import uvicorn

    from llmss_b200.serving.broker import connect
    from llmss_b200.serving.http import create_app

    args = get_args()
    app = create_app(lambda: connect(args.redis_host, args.redis_port))
    uvicorn.run(app, host=args.fastapi_host, port=args.fastapi_port)


if __name__ == "__main__":
    main()
