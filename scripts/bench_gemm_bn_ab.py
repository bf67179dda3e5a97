import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_gemm import bench
print("MAX_BN", os.environ.get("LLMSS_GEMM_MAX_BN"), flush=True)
for M in (256, 512, 2048, 8192):
    for N, K in ((4096, 2560), (3584, 4096), (4096, 4096), (12288, 4096), (28672, 4096), (4096, 20480)):
        if M > 512 and N * K > 60e6: continue
        bench(M, N, K)
