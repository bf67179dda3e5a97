"""One prefill-attention shape, a few launches: target for `ncu -k regex:attn_prefill`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llmss_b200.ops import cuda_impl as cu

nseq, L, hq, hkv, d = [int(x) for x in (sys.argv[1:6] if len(sys.argv) >= 6 else (4, 4096, 32, 32, 128))]
T = nseq * L
qkv = torch.randn(T, (hq + 2 * hkv) * d, device="cuda", dtype=torch.bfloat16)
out = torch.empty(T, hq * d, device="cuda", dtype=torch.bfloat16)
cu_seq = torch.arange(nseq + 1, dtype=torch.int32, device="cuda") * L
q, k, v = qkv[:, :hq * d], qkv[:, hq * d:(hq + hkv) * d], qkv[:, (hq + hkv) * d:]
for _ in range(5):
    cu.attn_prefill(q, k, v, out, cu_seq, nseq, L, hq, hkv, d, d ** -0.5)
torch.cuda.synchronize()
print("done")
