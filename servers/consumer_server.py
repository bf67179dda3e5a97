"""Consumer / tensor-parallel model server -- same flags as the reference (consumer_server.py:22-30).

    torchrun --nnodes 1 --nproc_per_node 4 servers/consumer_server.py --pretrained_model_path P \
        --redis_host 127.0.0.1 --redis_port 20000
`--model_type` (which the reference README passes but its parser rejects) is accepted and ignored.
"""
import os
import sys
from argparse import ArgumentParser

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def get_args():
    parser = ArgumentParser()
    consumer_group = parser.add_argument_group("consumer")
    consumer_group.add_argument("--pretrained_model_path", type=str, required=True)
    consumer_group.add_argument("--model_type", type=str, default=None)
    consumer_group.add_argument("--dtype", type=str, default="float16", choices=["float16", "bfloat16", "float32"])
    consumer_group.add_argument("--max_batch", type=int, default=64)
    consumer_group.add_argument("--max_seq_len", type=int, default=None)
    consumer_group.add_argument("--prefill_chunk", type=int, default=None,
                                help="feed prompts into the cache this many tokens per sequence and round (bounds activation memory)")
    consumer_group.add_argument("--prefix_cache", action="store_true",
                                help="prompts that start with the same tokens share the KV blocks of the common prefix")
    consumer_group.add_argument("--sliding_window", action="store_true",
                                help="past max_seq_len, keep generating over the most recent tokens (rotary-position models)")
    broker_group = parser.add_argument_group("broker")
    broker_group.add_argument("--redis_host", type=str, default="127.0.0.1")
    broker_group.add_argument("--redis_port", type=int, default=20000)
    return parser.parse_args()


def main():
    import torch

    from llmss_b200.engine.engine import Engine
    from llmss_b200.serving.broker import connect
    from llmss_b200.serving.consumer import ConsumerLoop
    from llmss_b200.utils.dist import initialize_torch_distributed

    args = get_args()
    process_group, rank, _ = initialize_torch_distributed()
    broker = connect(args.redis_host, args.redis_port) if rank == 0 else None
    tokenizer = None
    if rank == 0:
        from transformers import AutoTokenizer

        tokenizer = AutoTokenizer.from_pretrained(args.pretrained_model_path, padding_side="left", truncation_side="left")
    engine = Engine.from_pretrained(args.pretrained_model_path, dtype=getattr(torch, args.dtype),
                                    process_group=process_group, max_batch=args.max_batch, max_seq_len=args.max_seq_len,
                                    sliding_window=args.sliding_window, prefix_cache=args.prefix_cache,
                                    prefill_chunk=args.prefill_chunk)
    engine.comm.barrier()
    ConsumerLoop(engine, tokenizer, broker).run_forever()


if __name__ == "__main__":
    main()
