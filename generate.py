"""Batch generation CLI -- same flags and output as the reference `generate.py:21-32,192-196`.

    torchrun --nproc_per_node N generate.py --pretrained_model_path P --prompts "a" "b" \
        --max_new_tokens 20 [--is_greedy] [--temperature 1.0] [--top_p 0.95] [--top_k 50] [--use_cache]

Differences from the reference (all deliberate):
  * runs at WORLD_SIZE=1 and on CPU (Q2, Q4);                * sampling flags take effect (Q1);
  * `max_position_embeddings` is honoured when `n_positions` is absent (Q5);
  * prompts keep their own lengths -- no pad tokens enter the model (Q6);
  * `--use_cache` is accepted for compatibility; the paged KV cache is always on (Q12);
  * generation stops at the model's maximum length instead of a sliding window (Q7).
Extra flags: --dtype {float16,bfloat16,float32} (default float16 like the reference), --seed.
"""
import os
from argparse import ArgumentParser
from time import time

import torch

from llmss_b200.engine.engine import Engine
from llmss_b200.engine.runner import SamplingParams
from llmss_b200.parallel.comm import comm_for
from llmss_b200.utils.dist import initialize_torch_distributed


def get_args(argv=None):
    parser = ArgumentParser()
    parser.add_argument("--pretrained_model_path", type=str, required=True)
    parser.add_argument("--prompts", type=str, nargs="+", required=True)
    parser.add_argument("--max_new_tokens", type=int, default=20)
    parser.add_argument("--is_greedy", action="store_true")
    parser.add_argument("--temperature", type=float, default=1.0, help="If you don't wanna use this, set to 1.0")
    parser.add_argument("--top_p", type=float, default=0.95, help="If you don't wanna use this, set to 1.0")
    parser.add_argument("--top_k", type=int, default=50, help="If you don't wanna use this, set to 0")
    parser.add_argument("--use_cache", action="store_true")
    parser.add_argument("--dtype", type=str, default="float16", choices=["float16", "bfloat16", "float32"])
    parser.add_argument("--seed", type=int, default=0)
    parser.add_argument("--sliding_window", action="store_true",
                        help="keep generating past the model's context over the most recent tokens (rotary-position models)")
    return parser.parse_args(argv)


def main(argv=None):
    args = get_args(argv)
    assert args.max_new_tokens > 0, "Value of max_new_tokens should be over than 0."
    assert 0.0 < args.temperature <= 1.0, "Value of temperature is not valid."
    assert 0.0 < args.top_p <= 1.0, "Value of top_p is not valid."
    assert args.top_k >= 0, "Value of top_k is not valid."

    process_group, rank, world_size = initialize_torch_distributed()
    if rank == 0:
        start_time = time()

    from transformers import AutoTokenizer

    tokenizer = AutoTokenizer.from_pretrained(args.pretrained_model_path, padding_side="left", truncation_side="left")
    device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    comm = comm_for(process_group, device)
    comm.barrier()

    engine = Engine.from_pretrained(args.pretrained_model_path, dtype=getattr(torch, args.dtype), device=device,
                                    process_group=process_group, max_batch=max(1, len(args.prompts)),
                                    eos_token_id=None, sliding_window=args.sliding_window)
    max_len = engine.runner.max_seq_len

    # rank 0 tokenises; ids are shared with one broadcast (reference N5, generate.py:73-85)
    list_of_prompts = args.prompts
    if rank == 0:
        enc = tokenizer(list_of_prompts, return_attention_mask=False, truncation=True, max_length=max_len - 1)["input_ids"]
        payload = [list(map(int, e)) for e in enc]
    else:
        payload = None
    prompt_ids = comm.broadcast_object(payload, src=0)

    params = SamplingParams(max_new_tokens=args.max_new_tokens, is_greedy=args.is_greedy, temperature=args.temperature,
                            top_p=args.top_p, top_k=args.top_k, seed=args.seed)
    # --use_cache off: the reference recomputes the whole sequence every step (generate.py:147-190); same tokens, O(S^2)
    outputs = engine.generate(prompt_ids, params) if (args.use_cache or os.environ.get("LLMSS_ALWAYS_CACHE", "0") == "1") \
        else engine.generate_no_cache(prompt_ids, params)

    if rank == 0:
        end_time = time()
        print(f"elapsed time: {end_time - start_time}")
        print(f"prompts: {list_of_prompts}")
        print(f"continuations: {[tokenizer.decode(o) for o in outputs]}")
    return outputs


if __name__ == "__main__":
    main()
