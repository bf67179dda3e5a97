"""Grid-size sensitivity of the stream-K GEMM (cold weights): does a tile-aligned grid beat 148 CTAs + fix-up?"""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llmss_b200.ops import cuda_impl as cu
from scripts.bench_gemm import timeit

for (M, N, K, grids) in ((64, 28672, 4096, (0, 112, 148)), (64, 14336, 4096, (0, 112, 148)), (64, 12288, 4096, (0, 96, 148)),
                         (64, 16384, 4096, (0, 128, 148)), (64, 50432, 4096, (0, 132, 148)), (64, 11008, 4096, (0, 86, 148)),
                         (16, 28672, 4096, (0, 112, 148))):
    nv = max(2, int(400e6 // (N * K * 2)) + 1)
    ws = [torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / math.sqrt(K) for _ in range(nv)]
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    bias = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    res = []
    for g in grids:
        t = timeit(lambda i: cu.linear(x, ws[i], out=out, bias=bias, act=1, max_ctas=g), nv)
        res.append(f"grid {g or 'auto':>4}: {t[1]:6.1f} us")
    print(f"M={M} N={N} K={K} tiles={(N + 127) // 128} | " + " | ".join(res), flush=True)
    del ws
