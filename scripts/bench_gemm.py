"""Device-timed GEMM micro-benchmark: tcgen05 stream-K kernel vs cuBLAS (torch F.linear).

Weights are cycled through a pool larger than L2 (126 MB) so every timed launch streams its
weights from HBM; CUDA events on the launching stream, best-of and median reported.
usage: python scripts/bench_gemm.py [--json out.json]
"""
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from llmss_b200.ops import cuda_impl as cu

PEAK_BW = 6545.6e9
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    PEAK_BW = PEAK["hbm_gbs"] * 1e9
    PEAK_FLOPS = PEAK["bf16_tflops"] * 1e12
except Exception:
    PEAK_FLOPS = 1654.1e12


def timeit(fn, n_variants, iters=10, warmup=3, per_graph=None):
    """Per-launch device time with launch gaps removed: `per_graph` launches (cycling through
    the weight variants, > L2 in total) are captured in one CUDA graph; events bracket replays."""
    per_graph = per_graph or max(8, n_variants)
    for i in range(warmup):
        fn(i % n_variants)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(per_graph):
            fn(i % n_variants)
    g.replay()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record()
        g.replay()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) * 1e3 / per_graph for a, b in evs)  # us per launch
    return ts[0], ts[len(ts) // 2]


def bench(M, N, K, dtype=torch.bfloat16):
    wbytes = N * K * 2
    nvar = max(2, min(64, int(400e6 // wbytes) + 1))  # > 2 x L2 worth of distinct weights
    ws = [torch.randn(N, K, device="cuda", dtype=dtype) / math.sqrt(K) for _ in range(nvar)]
    x = torch.randn(M, K, device="cuda", dtype=dtype)
    out = torch.empty(M, N, device="cuda", dtype=dtype)
    p = cu.gemm_plan(M, N, K)
    ours = timeit(lambda i: cu.linear(x, ws[i], out=out), nvar)
    blas_out = torch.empty(M, N, device="cuda", dtype=dtype)
    blas = timeit(lambda i: torch.mm(x, ws[i].t(), out=blas_out), nvar)
    bytes_ = wbytes + M * K * 2 + M * N * 2
    flops = 2.0 * M * N * K
    roof = max(bytes_ / PEAK_BW, flops / PEAK_FLOPS) * 1e6
    r = dict(M=M, N=N, K=K, bn=p.bn, grid=p.grid, ours_us=ours[1], ours_best_us=ours[0], cublas_us=blas[1],
             cublas_best_us=blas[0], roofline_us=roof, frac_of_roofline=roof / ours[1], speedup_vs_cublas=blas[1] / ours[1])
    print(f"M={M:5d} N={N:6d} K={K:6d} bn={p.bn:3d} grid={p.grid:3d} | ours {ours[1]:8.1f} us (best {ours[0]:7.1f}) | "
          f"cuBLAS {blas[1]:8.1f} us | roofline {roof:7.1f} us -> {100 * roof / ours[1]:5.1f}% of measured roofline, "
          f"{blas[1] / ours[1]:.2f}x cuBLAS", flush=True)
    return r


if __name__ == "__main__":
    shapes = []
    for M in ([int(a) for a in os.environ["BENCH_MS"].split(",")] if "BENCH_MS" in os.environ else (1, 16, 64, 256)):
        shapes += [(M, 28672, 4096), (M, 4096, 20480), (M, 12288, 4096), (M, 4096, 4096), (M, 16384, 4096),
                   (M, 4096, 16384), (M, 3584, 4096), (M, 4096, 2560), (M, 50432, 4096)]
    if "BENCH_MS" not in os.environ:
        shapes += [(2048, 12288, 4096), (8192, 4096, 4096), (8192, 16384, 4096), (8192, 4096, 16384)]
    res = [bench(*s) for s in shapes]
    if "--json" in sys.argv:
        json.dump(res, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)
