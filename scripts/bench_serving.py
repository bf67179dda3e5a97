"""Serving benchmark through the real front ends (BASELINE.json configs 3 and 4), one process per GPU.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 scripts/bench_serving.py --model llama-2-13b \
        --frontend pubsub --batch 64 --prompt-len 1024 --gen 256

All TP ranks run the consumer loop (`serving/consumer.py`, the counterpart of the reference's
`consumer_server.py:75-173`).  Rank 0 spawns ONE child process that hosts what the reference runs as separate
processes: the broker (RESP wire protocol over TCP, the `redis-server --port 20000` of the PoC README), the
front end (`grpc`: the gRPC service; `pubsub`: clients that LPUSH to `pqueue` / BRPOP their reply list exactly like
`producer_server.py:45-55`; `http`: the FastAPI producer), and `--batch` concurrent clients, each sending one
request with a `--prompt-len`-token prompt and `--gen` new tokens.  Everything between client and engine is real
(TCP, JSON, protobuf); only the tokenizer is a byte-level stand-in (no tokenizer files offline: one token per
byte, so a 1024-character prompt is a 1024-token prompt).

Reported (one JSON line from rank 0): wall-clock tokens/s from first send to last answer, p50 / p99 TTFT and TPOT
as measured by the engine per request (returned in every response), and the client-side p50 latency.  The
`Engine.step()` TPOT at the same batch (bench.py) is the number to hold it against.
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


class ByteTokenizer:
    """One token per byte (shifted past the special ids); stands in for the HF tokenizer files."""

    def __init__(self, vocab: int):
        self.vocab = vocab

    def __call__(self, text, return_attention_mask=False, truncation=True, max_length=None, **_):
        ids = [3 + (b % (self.vocab - 8)) for b in text.encode("latin-1")]
        if max_length:
            ids = ids[-max_length:]
        return {"input_ids": ids}

    def decode(self, ids):
        return bytes(max(0, i - 3) % 256 for i in ids).decode("latin-1")


def client_process(args):
    """Child process of rank 0: broker server + front end + `batch` concurrent clients (what the reference runs as
    `redis-server`, `producer_server.py` and curl).  Writes its measurements to args.result_file."""
    import random

    from llmss_b200.serving.broker import PRODUCER_QUEUE, BrokerClient, BrokerServer
    from llmss_b200.serving.protocol import GenerateRequest, GenerateResponse

    server = BrokerServer("127.0.0.1", args.port).start()
    rnd = random.Random(0)
    lat, resp = [None] * args.batch, [None] * args.batch
    kw = dict(max_new_tokens=args.gen, is_greedy=not args.sampling, temperature=0.8, top_p=0.95, top_k=50)

    def prompt():
        return "".join(chr(33 + rnd.randrange(90)) for _ in range(args.prompt_len))

    grpc_client = front = None
    if args.frontend == "grpc":
        from llmss_b200.serving.grpc_service import Client, serve

        front, gport = serve(lambda: BrokerClient("127.0.0.1", args.port), "127.0.0.1", 0, max_workers=args.batch + 8)
        grpc_client = Client(f"127.0.0.1:{gport}")
        assert grpc_client.health() == "ok"
    elif args.frontend == "http":
        import uvicorn

        from llmss_b200.serving.http import create_app

        app = create_app(lambda: BrokerClient("127.0.0.1", args.port))
        front = uvicorn.Server(uvicorn.Config(app, host="127.0.0.1", port=args.port + 1, log_level="error"))
        threading.Thread(target=front.run, daemon=True).start()
        time.sleep(1.0)

    def one(i, text, store=True):
        t0 = time.perf_counter()
        if args.frontend == "grpc":
            r = grpc_client.generate(text, **kw)
            out = dict(num_tokens=r.num_tokens, ttft_ms=r.ttft_ms, tpot_ms=r.tpot_ms)
        elif args.frontend == "http":
            import requests

            r = requests.post(f"http://127.0.0.1:{args.port + 1}/generate", json=dict(prompt=text, **kw), timeout=600)
            r.raise_for_status()
            out = dict(num_tokens=args.gen, ttft_ms=None, tpot_ms=None)
        else:  # pub/sub: the producer's own wire protocol (LPUSH pqueue / BRPOP reply list)
            c = BrokerClient("127.0.0.1", args.port)
            req = GenerateRequest(prompt=text, **kw).with_routing()
            c.lpush(PRODUCER_QUEUE, req.to_json())
            g = GenerateResponse.from_json(c.brpop(req.reply_to, timeout=600)[1])
            c.close()
            assert g.error is None, g.error
            out = dict(num_tokens=g.num_tokens, ttft_ms=g.ttft_ms, tpot_ms=g.tpot_ms)
        if store:
            lat[i], resp[i] = time.perf_counter() - t0, out

    # warm-up: the whole batch once (captures the decode graph for this bucket, tunes the attention split)
    ws = [threading.Thread(target=one, args=(i, prompt(), False)) for i in range(args.batch)]
    [t.start() for t in ws]
    [t.join() for t in ws]
    texts = [prompt() for _ in range(args.batch)]
    t0 = time.perf_counter()
    ts = [threading.Thread(target=one, args=(i, texts[i])) for i in range(args.batch)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    wall = time.perf_counter() - t0
    ntok = sum(r["num_tokens"] for r in resp)
    ttft = [r["ttft_ms"] for r in resp if r["ttft_ms"]]
    tpot = [r["tpot_ms"] for r in resp if r["tpot_ms"]]
    q = lambda xs, p: (sorted(xs)[min(len(xs) - 1, int(p * len(xs)))] if xs else None)  # noqa: E731
    json.dump({"value": round(ntok / wall, 1), "wall_s": round(wall, 3), "tokens": ntok,
               "ttft_p50_ms": q(ttft, 0.5), "ttft_p99_ms": q(ttft, 0.99), "tpot_p50_ms": q(tpot, 0.5),
               "tpot_p99_ms": q(tpot, 0.99), "client_latency_p50_s": round(statistics.median(lat), 3)},
              open(args.result_file, "w"))
    c = BrokerClient("127.0.0.1", args.port)
    c.lpush(PRODUCER_QUEUE, GenerateRequest(prompt="__shutdown__").to_json())
    time.sleep(1.0)   # let the consumer pop it before the broker goes away
    if front is not None and hasattr(front, "stop"):
        front.stop(0)
    server.stop()


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-2-7b")
    ap.add_argument("--frontend", choices=["grpc", "pubsub", "http"], default="grpc")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--prompt-len", type=int, default=1024)
    ap.add_argument("--gen", type=int, default=256)
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--device", default="cuda")
    ap.add_argument("--port", type=int, default=20000)
    ap.add_argument("--sampling", action="store_true", help="top-k / top-p requests instead of greedy")
    ap.add_argument("--ckpt-dir", default=os.environ.get("LLMSS_BENCH_CKPT", "/tmp/llmss_b200_bench"))
    ap.add_argument("--json", default=None)
    ap.add_argument("--role", choices=["server", "client"], default="server")
    ap.add_argument("--result-file", default=None)
    return ap.parse_args()


def main():
    args = parse()
    if args.role == "client":
        return client_process(args)

    import subprocess
    import tempfile

    from llmss_b200.engine.engine import Engine
    from llmss_b200.serving.broker import BrokerClient
    from llmss_b200.serving.consumer import ConsumerLoop
    from llmss_b200.utils.checkpoint import write_random_checkpoint
    from llmss_b200.utils.dist import initialize_torch_distributed

    pg, rank, world = initialize_torch_distributed(timeout_s=600)
    local = int(os.environ.get("LOCAL_RANK", rank))
    cuda = args.device == "cuda"
    dev = torch.device("cuda", local) if cuda else torch.device("cpu")
    path = os.path.join(args.ckpt_dir, args.model)
    if rank == 0:
        write_random_checkpoint(path, args.model, dtype=torch.float16 if cuda else torch.float32, device=args.device)
    if world > 1:
        dist.barrier(device_ids=[local]) if cuda else dist.barrier()
    dtype = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}[args.dtype]
    eng = Engine.from_pretrained(path, dtype=dtype, device=dev, process_group=pg, max_batch=args.batch,
                                 max_seq_len=args.prompt_len + args.gen + 16, max_prefill_tokens=8192,
                                 symmetric_bytes=128 << 20, lamport_bytes=64 << 20)
    tok = ByteTokenizer(eng.core.spec.vocab)
    child = broker = None
    result_file = None
    if rank == 0:
        result_file = tempfile.mktemp(suffix=".json")
        child = subprocess.Popen([sys.executable, os.path.abspath(__file__), "--role", "client", "--result-file", result_file]
                                 + [a for a in sys.argv[1:]])
        for _ in range(600):   # the child owns the broker: wait until it listens
            try:
                broker = BrokerClient("127.0.0.1", args.port)
                broker.ping()
                break
            except OSError:
                broker = None
                time.sleep(0.1)
        assert broker is not None, "broker did not come up"
    loop = ConsumerLoop(eng, tok, broker, max_new_per_step=args.batch, idle_block_s=0.01)
    loop.run_forever()
    if rank == 0:
        child.wait(timeout=60)
        result = {"metric": "serving_tokens_per_s", "unit": "tokens/s", "n_gpus": world, "frontend": args.frontend,
                  "dtype": args.dtype, "data": "synthetic (byte-level tokenizer stand-in)",
                  "config": {"model": f"{args.model} (random-init)", "batch": args.batch, "prompt_len": args.prompt_len,
                             "gen": args.gen, "parallelism": f"tp{world}", "greedy": not args.sampling}}
        result.update(json.load(open(result_file)))
        from llmss_b200.utils.metrics import METRICS

        ds, ls = METRICS.quantile("decode_step_seconds", 0.5), METRICS.quantile("consumer_step_seconds", 0.5)
        result.update({"engine_decode_step_p50_ms": None if ds is None else round(1e3 * ds, 4),
                       "consumer_decode_iteration_p50_ms": None if ls is None else round(1e3 * ls, 4),
                       "note": "tpot_* is per request, (last - first token) / (n - 1): it includes the prefill steps of "
                               "requests admitted later (continuous batching); engine_decode_step is Engine.step() for a "
                               "decode step, consumer_decode_iteration the whole serving-loop iteration around it",
                       "decode_steps_total": eng.runner.stats["decode_steps"], "plan_bytes_shared_total": loop.bytes_shared,
                       "heap": {"mode": getattr(eng.comm.symm, "mode", None),
                                "nvls": bool(getattr(eng.comm.symm, "has_multicast", False))}})
        print(json.dumps(result), flush=True)
        if args.json:
            json.dump(result, open(args.json, "w"), indent=1)
    if world > 1:
        dist.barrier(device_ids=[local]) if cuda else dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
