"""Epilogue-flavour sensitivity of the CTA-pair GEMM kernel against the single-CTA kernels (same timing method as
scripts/bench_gemm_pair.py): the two biggest prefill GEMMs of the model families -- GPT-J's fused QKV + fc_in (bias +
GELU from feature 12288 on) and Llama-7B's gate/up (SwiGLU on interleaved rows) -- and their plain forms."""
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llmss_b200.ops import _lib
from llmss_b200.ops import cuda_impl as cu
from llmss_b200.ops import reference as R
from bench_gemm import timeit  # noqa: E402

lib = _lib.load()
rows = []
for M in (8192, 256):
    for name, N, K, kw in (("gptj qkv+fc_in bias+gelu", 28672, 4096, dict(act=R.ACT_GELU_TANH, act_start=12288, bias=True)),
                           ("plain", 28672, 4096, dict()),
                           ("llama gate/up swiglu", 22016, 4096, dict(act=R.ACT_SILU, gated=True)),
                           ("plain", 22016, 4096, dict()),
                           ("gptj out+down residual", 4096, 20480, dict(residual=True, bias=True))):
        nvar = 3
        ws = [torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / math.sqrt(K) for _ in range(nvar)]
        x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
        kw = dict(kw)
        if kw.pop("bias", False):
            kw["bias"] = torch.randn(N, device="cuda", dtype=torch.bfloat16)
        if kw.pop("residual", False):
            kw["residual"] = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
        out = torch.empty(M, N // 2 if kw.get("gated") else N, device="cuda", dtype=torch.bfloat16)
        t = {}
        for label, mode in (("single", 0), ("pair", -1)):
            lib.llmss_gemm_set_pair(mode)
            t[label] = timeit(lambda i: cu.linear(x, ws[i], out=out, **kw), nvar, iters=6)[1]
        lib.llmss_gemm_set_pair(-1)
        rows.append(dict(M=M, N=N, K=K, flavour=name, single_us=t["single"], pair_us=t["pair"]))
        print(f"M={M:5d} N={N:6d} K={K:6d} {name:26s} single-CTA {t['single']:8.1f} us | pair {t['pair']:8.1f} us | {t['pair'] / t['single']:.2f}x", flush=True)
if "--json" in sys.argv:
    json.dump(rows, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)
