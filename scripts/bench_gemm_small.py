"""TP=8-sized decode GEMMs (per-rank shapes) vs cuBLAS; A/B of the split-K heuristic via LLMSS_GEMM_MIN_KB_PER_CTA."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_gemm import bench  # noqa: E402

if __name__ == "__main__":
    print("LLMSS_GEMM_MIN_KB_PER_CTA =", os.environ.get("LLMSS_GEMM_MIN_KB_PER_CTA", "(default 8)"), flush=True)
    for M in (16, 64, 128):
        for N, K in ((4096, 512), (4096, 1344), (5120, 1728), (4096, 2560), (3584, 4096), (1536, 4096), (2752, 4096), (6304, 4096)):
            bench(M, N, K)
