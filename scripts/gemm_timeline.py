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


def run_pair(M, N, K, split, dbg_mode=0):
    """CTA-pair kernel (cta_group::2): stamps 0 entry, 1 prologue, 3 first stage landed (issuer), 4 last MMA issued,
    5 accumulator ready, 2 partials parked + cluster barrier, 6 epilogue done, 7 exit."""
    lib = _lib.load()
    w = [torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / math.sqrt(K) for _ in range(3)]
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    tiles = (-(-N // 256)) * (-(-M // 256))
    rounds = -(-tiles // 74)
    grid = (tiles if split > 1 else -(-tiles // rounds)) * 2 * split
    os.environ["LLMSS_GEMM_PAIR_DEBUG"] = str(dbg_mode)
    lib.llmss_gemm_set_pair(split)
    for i in range(3):
        cu.linear(x, w[i], out=out)
    dbg = torch.zeros(grid * 8, dtype=torch.int64, device="cuda")
    lib.llmss_gemm_set_debug(C.c_void_p(dbg.data_ptr()))
    cu.linear(x, w[0], out=out)
    torch.cuda.synchronize()
    lib.llmss_gemm_set_debug(None)
    lib.llmss_gemm_set_pair(0)
    os.environ["LLMSS_GEMM_PAIR_DEBUG"] = "0"
    t = dbg.view(grid, 8).cpu().double()
    rel = (t - t[:, 0].min()) / 1e3
    what = {0: "normal", 2: "no MMAs (loads only)", 3: "no loads (MMAs only)"}[dbg_mode]
    print(f"PAIR M={M} N={N} K={K} pairs/cluster={split} grid={grid} [{what}]: kernel span {rel[:, 7].max():.1f} us, "
          f"k-blocks per CTA {K // 64 // split}")
    for i, n in [(0, "entry"), (1, "prologue done"), (3, "first stage landed"), (4, "last MMA issued"), (5, "accumulator ready"),
                 (2, "parked + barrier"), (6, "epilogue done"), (7, "exit")]:
        col = rel[:, i][t[:, i] > 0]
        if len(col):
            print(f"   {n:20s} min {col.min():7.2f}  median {col.median():7.2f}  max {col.max():7.2f} us")
    lead = t[:, 4] > 0
    per = ((t[:, 4] - t[:, 3])[lead] / 1e3 / max(1, K // 64 // split - 1))
    print(f"   issuer: first stage -> last MMA issued = {((t[:, 4] - t[:, 3])[lead] / 1e3).median():.2f} us median "
          f"= {per.median() * 1e3:.0f} ns per k-block")


if len(sys.argv) > 1 and sys.argv[1] == "pair":
    for mode in (0, 2, 3):
        run_pair(256, 4096, 4096, 1, mode)
    run_pair(256, 4096, 4096, 2, 0)
    for mode in (0, 2, 3):
        run_pair(8192, 4096, 4096, 1, mode)
    sys.exit(0)
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
