"""Launch one GEMM shape a few times (for ncu).  usage: prof_gemm.py M N K [iters]"""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llmss_b200.ops import cuda_impl as cu

M, N, K = (int(a) for a in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 6
ws = [torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / math.sqrt(K) for _ in range(3)]
x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
kw = {}
if os.environ.get("EPI") == "gelu":  # the fused QKV+up GEMM of the GPT-J block: bias + GELU on the MLP columns
    kw = dict(bias=torch.randn(N, device="cuda", dtype=torch.bfloat16), act=1, act_start=12288 if N > 12288 else 0)
for i in range(iters):
    cu.linear(x, ws[i % 3], out=out, **kw)
torch.cuda.synchronize()
print("ok", cu.gemm_plan(M, N, K).grid)
