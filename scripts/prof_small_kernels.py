"""The small kernels of a decode step at flagship shapes (GPT-J-6B, batch 64), a few launches each: ncu target."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llmss_b200.ops import cuda_impl as cu
from llmss_b200.ops import reference as R

B, H, V, hq, d = 64, 4096, 50400, 16, 256
dev, dt = "cuda", torch.bfloat16
x = torch.randn(B, H, device=dev, dtype=dt)
res = torch.randn(B, H, device=dev, dtype=dt)
w, b = torch.randn(H, device=dev, dtype=dt), torch.randn(H, device=dev, dtype=dt)
wte = torch.randn(V, H, device=dev, dtype=dt)
tok = torch.randint(0, V, (B,), device=dev, dtype=torch.int32)
width = 3 * hq * d
qkv = torch.randn(B, width, device=dev, dtype=dt)
pos = torch.full((B,), 1040, device=dev, dtype=torch.int32)
nblk, bs = 64 * 18, 64
kc = torch.zeros(nblk, hq, bs, d, device=dev, dtype=dt)
vc = torch.zeros_like(kc)
slots = (torch.arange(B, device=dev, dtype=torch.int32) * 17 * bs + 16)
logits = torch.randn(B, V + 32, device=dev, dtype=torch.float32) * 3
greedy = torch.tensor([1, 0] * (B // 2), device=dev, dtype=torch.int32)
temp = torch.full((B,), 0.8, device=dev)
topk = torch.full((B,), 50, device=dev, dtype=torch.int32)
topp = torch.full((B,), 0.95, device=dev)
rng = torch.zeros(B, 2, device=dev, dtype=torch.int32)
for _ in range(3):
    cu.embed(tok, wte)
    cu.norm_residual(x, res, w, b, 1e-5, R.NORM_LAYER)
    cu.rope_kv_write(qkv, pos, slots, kc, vc, hq, hq, d, 64, R.ROPE_GPTJ, 10000.0, 2 * hq * d, 0, hq * d)
    cu.sample(logits, greedy, temp, topk, topp, rng, vocab=V)
    rec = cu.vp_summary(logits[:, :6304].contiguous(), 6300, 0, greedy, topk)
    cu.vp_merge_sample(torch.cat([rec] * 8, dim=1).contiguous(), 8, V, greedy, temp, topk, topp, rng)
torch.cuda.synchronize()
print("ok")
