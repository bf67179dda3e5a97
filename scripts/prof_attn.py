"""Launch the decode-attention kernel a few times for one shape (for ncu). usage: prof_attn.py B hq hkv d ctx [splits]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llmss_b200.ops import cuda_impl as cu

B, hq, hkv, d, ctx = (int(a) for a in sys.argv[1:6])
bs = int(os.environ.get("BS", "16"))
if os.environ.get("TC_ALL"):
    cu._TC_MIN_GROUP = 1
maxb = (ctx + bs - 1) // bs
nb = 1 + B * maxb
kc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=torch.bfloat16)
vc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=torch.bfloat16)
bt = (torch.arange(B * maxb, device="cuda", dtype=torch.int32) + 1).view(B, maxb)
sl = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
q = torch.randn(B, hq * d, device="cuda", dtype=torch.bfloat16)
out = torch.empty_like(q)
sp = int(sys.argv[6]) if len(sys.argv) > 6 else cu.decode_splits(B, hq, hkv, d, ctx, torch.bfloat16, bs)
for _ in range(5):
    cu.attn_decode(q, kc, vc, bt, sl, out, hq, hkv, d, d ** -0.5, sp)
torch.cuda.synchronize()
print("ok splits", sp)
