"""ncu target: the CTA-pair GEMM kernel on a decode-side shape (one tile per pair, exposed epilogue) and a prefill shape
(7 tiles per pair, epilogue under the next main loop).  Launch order: A B A B A B (ncu: -s 4 -c 2)."""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llmss_b200.ops import cuda_impl as cu

shapes = [(256, 16384, 4096), (8192, 4096, 4096)]
ops = []
for M, N, K in shapes:
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / math.sqrt(K)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    ops.append((x, w, out))
for _ in range(3):
    for x, w, out in ops:
        cu.linear(x, w, out=out)
torch.cuda.synchronize()
print("ok")
