"""Epilogue-flavour sensitivity of the decode GEMM (cold weights): plain vs bias vs bias + partial GELU."""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llmss_b200.ops import cuda_impl as cu
from scripts.bench_gemm import timeit

M, N, K = 64, 28672, 4096
if len(sys.argv) > 3:
    M, N, K = (int(a) for a in sys.argv[1:4])
ws = [torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / math.sqrt(K) for _ in range(3)]
x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
bias = torch.randn(N, device="cuda", dtype=torch.bfloat16)
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
res = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
for name, kw in (("plain", {}), ("bias", dict(bias=bias)), ("bias+gelu[12288:]", dict(bias=bias, act=1, act_start=12288)),
                 ("bias+gelu[all]", dict(bias=bias, act=1)), ("bias+residual", dict(bias=bias, residual=res))):
    if kw.get("act_start", 0) >= N:
        continue
    t = timeit(lambda i: cu.linear(x, ws[i], out=out, **kw), 3)
    print(f"M={M} N={N} K={K} {name:20s} {t[1]:7.1f} us (best {t[0]:.1f})", flush=True)
for mc in (0, 112, 128, 140, 148):
    t = timeit(lambda i: cu.linear(x, ws[i], out=out, max_ctas=mc), 3)
    print(f"max_ctas={mc:3d} plain {t[1]:7.1f} us", flush=True)
