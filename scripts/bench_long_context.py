"""Long-context numbers: time to first token and time per output token for 8k / 16k-token prompts.

    python scripts/bench_long_context.py [--model llama-2-7b] [--json out.json]

One process, one GPU, random-init weights of the named architecture (RoPE models have no position table, so the
engine's `max_seq_len` is the only limit).  The reference cannot run this at all: its attention materialises an
fp32 [B, H, S, T] score tensor (gptj_modeling.py:141-167) -- 32 heads x 16k x 16k x 4 B = 34 GB for ONE sequence --
and its KV cache is re-allocated by `torch.cat` every step.  Here: tcgen05 flash prefill (whole prompt, or in chunks
over the paged cache) and the paged tcgen05 decode kernel.  TTFT = host wall clock around the prefill step(s) including
the D2H read of the first token; TPOT = median of 32 `Engine.step()` decode steps after it.
"""
import argparse
import json
import os
import statistics
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llmss_b200.engine.engine import Engine
from llmss_b200.engine.runner import SamplingParams
from llmss_b200.utils.checkpoint import write_random_checkpoint


def run(path, n_seqs, prompt_len, chunk, vocab, n_decode=32):
    eng = Engine.from_pretrained(path, dtype=torch.bfloat16, device="cuda", max_batch=max(1, n_seqs),
                                 max_seq_len=prompt_len + n_decode + 64, max_prefill_tokens=max(8192, prompt_len),
                                 prefill_chunk=chunk)
    g = torch.Generator().manual_seed(prompt_len + n_seqs)
    sp = SamplingParams(max_new_tokens=n_decode + 1, is_greedy=True)
    best = None
    for rep in range(2):   # first pass warms the kernels' one-time set-up (tensor maps, attributes); second is reported
        prompts = [torch.randint(3, vocab - 1, (prompt_len,), generator=g).tolist() for _ in range(n_seqs)]
        for p in prompts:
            eng.add_request(p, sp)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        first = 0
        while first < n_seqs:
            first += sum(1 for o in eng.step())
        torch.cuda.synchronize()
        ttft = time.perf_counter() - t0
        steps = []
        while eng.has_work():
            t1 = time.perf_counter()
            eng.step()
            steps.append(time.perf_counter() - t1)
        best = dict(n_seqs=n_seqs, prompt_len=prompt_len, prefill_chunk=chunk, ttft_ms=round(1e3 * ttft, 1),
                    prefill_tokens_per_s=round(n_seqs * prompt_len / ttft),
                    tpot_ms=round(1e3 * statistics.median(steps[2:]), 3), decode_steps=len(steps))
    del eng
    torch.cuda.empty_cache()
    print(json.dumps(best), flush=True)
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-2-7b")
    ap.add_argument("--ckpt-dir", default=os.environ.get("LLMSS_BENCH_CKPT", "/tmp/llmss_b200_bench"))
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    from llmss_b200.utils.checkpoint import PRESETS

    cfg = dict(PRESETS[args.model])
    for k in ("max_position_embeddings", "n_positions"):   # same weights, longer declared context (RoPE: no table)
        if k in cfg:
            cfg[k] = 32768
    path = os.path.join(args.ckpt_dir, args.model + "-32k")
    write_random_checkpoint(path, cfg, dtype=torch.float16, device="cuda")
    from transformers import AutoConfig

    vocab = AutoConfig.from_pretrained(path).vocab_size
    rows = [run(path, 1, 8192, None, vocab), run(path, 1, 8192, 2048, vocab), run(path, 1, 16384, None, vocab),
            run(path, 8, 8192, None, vocab)]
    if args.json:
        json.dump({"model": f"{args.model} (random-init)", "dtype": "bf16", "gpus": 1, "rows": rows}, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
