"""Detailed diagnosis of the tcgen05 GEMM on a few shapes (prints error structure)."""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from llmss_b200.ops import cuda_impl as cu
from llmss_b200.ops import reference as R


def diag(M, N, K, dtype=torch.bfloat16, pattern="randn"):
    torch.manual_seed(0)
    if pattern == "ones":
        x = torch.ones(M, K, device="cuda", dtype=dtype)
        w = torch.ones(N, K, device="cuda", dtype=dtype) / K
    elif pattern == "eye":  # w row n selects x column (n % K): out[m, n] = x[m, n % K]
        x = torch.randn(M, K, device="cuda", dtype=dtype)
        w = torch.zeros(N, K, device="cuda", dtype=dtype)
        w[torch.arange(N), torch.arange(N) % K] = 1
    else:
        x = torch.randn(M, K, device="cuda", dtype=dtype)
        w = torch.randn(N, K, device="cuda", dtype=dtype) / math.sqrt(K)
    p = cu.gemm_plan(M, N, K)
    y = cu.linear(x, w)
    torch.cuda.synchronize()
    ref = R.linear(x, w, out_dtype=torch.float32)
    err = (y.float() - ref).abs()
    bad = err > (3e-2 + 2e-2 * ref.abs())
    print(f"[{pattern}] M={M} N={N} K={K} bn={p.bn} tiles={p.tiles} kb={p.kb_total} grid={p.grid}: "
          f"max_err={err.max().item():.4g} bad={int(bad.sum())}/{bad.numel()}", flush=True)
    if bad.any():
        rows = bad.any(1).nonzero().flatten().tolist()
        cols = bad.any(0).nonzero().flatten().tolist()
        print("   bad rows:", rows[:16], "... bad cols:", cols[:24], "(", len(rows), "rows", len(cols), "cols )")
        print("   got[0,:8] ", y[0, :8].float().tolist())
        print("   ref[0,:8] ", ref[0, :8].tolist())
        ratio = (y.float() / ref)[~torch.isnan(y.float() / ref)]
        print("   median got/ref:", ratio.median().item())
    return not bad.any()


ok = True
for pat in ("ones", "eye", "randn"):
    for shape in [(16, 128, 64), (16, 128, 256), (64, 256, 512), (64, 4096, 4096), (256, 1024, 1024), (1, 4096, 4096)]:
        ok &= diag(*shape, pattern=pat)
for dt in (torch.float16,):
    ok &= diag(32, 512, 512, dtype=dt)
print("GEMM_DEBUG", "OK" if ok else "FAILED")
sys.exit(0 if ok else 1)
