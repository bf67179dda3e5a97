"""Per-CTA phase timeline of the GEMM kernels (globaltimer stamps, see llmss_gemm_set_debug)."""
import ctypes as C
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llmss_b200.ops import _lib
from llmss_b200.ops import cuda_impl as cu

NAMES = ["entry", "prologue", "fixupDecided", "firstFull", "lastMMA", "accReady(last seg)", "epiDone", "exit"]


def run(M, N, K, **kw):
    lib = _lib.load()
    w = [torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / math.sqrt(K) for _ in range(3)]
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    p = cu.gemm_plan(M, N, K)
    if kw.pop("bias", False):
        kw["bias"] = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    for i in range(3):
        cu.linear(x, w[i], out=out, **kw)
    dbg = torch.zeros(p.grid * 8, dtype=torch.int64, device="cuda")
    lib.llmss_gemm_set_debug(C.c_void_p(dbg.data_ptr()))
    cu.linear(x, w[0], out=out, **kw)
    torch.cuda.synchronize()
    lib.llmss_gemm_set_debug(None)
    t = dbg.view(p.grid, 8).cpu().double()
    t0 = t[:, 0].min()
    rel = (t - t0) / 1e3
    print(f"M={M} N={N} K={K} {sorted(kw)} bn={p.bn} split={p.split} grid={p.grid}: kernel span {rel[:, 7].max():.1f} us")
    for i, n in enumerate(NAMES):
        col = rel[:, i][t[:, i] > 0]
        if len(col):
            print(f"   {n:18s} min {col.min():7.2f}  median {col.median():7.2f}  max {col.max():7.2f} us")


    if p.split == 0:
        d1 = (t[:, 2] - t[:, 5])[t[:, 2] > 0] / 1e3
        d2 = (t[:, 6] - t[:, 2])[t[:, 2] > 0] / 1e3
        d3 = (t[:, 6] - t[:, 5]) / 1e3
        print(f"   last segment: dump+fence+ticket median {d1.median():.2f} max {d1.max():.2f} us | after decision -> done "
              f"median {d2.median():.2f} max {d2.max():.2f} us | total median {d3.median():.2f} max {d3.max():.2f}")


if len(sys.argv) > 1 and sys.argv[1] == "small":
    run(64, 3584, 4096, bias=True, act=1, act_start=1536)
    run(64, 4096, 2560, bias=True)
    run(64, 4096, 4096)
    sys.exit(0)
if len(sys.argv) > 1 and sys.argv[1] == "tp8":
    run(64, 4096, 512)
    run(64, 4096, 2560)
    run(64, 3584, 4096, bias=True, act=1, act_start=1536)
    sys.exit(0)
if len(sys.argv) > 1 and sys.argv[1] == "epi":
    run(64, 28672, 4096)
    run(64, 28672, 4096, bias=True, act=1, act_start=12288)
    sys.exit(0)
for shape in [(64, 4096, 2560), (16, 4096, 2560), (64, 28672, 4096), (64, 4096, 20480), (8192, 4096, 4096), (2048, 12288, 4096)]:
    run(*shape)
