"""Device-timed causal prefill attention: own tcgen05 kernel vs torch SDPA (library)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from llmss_b200.ops import cuda_impl as cu


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / iters


def bench(nseq, L, hq, hkv, d, dtype=torch.bfloat16):
    T = nseq * L
    qkv = torch.randn(T, (hq + 2 * hkv) * d, device="cuda", dtype=dtype)
    out = torch.empty(T, hq * d, device="cuda", dtype=dtype)
    cu_seq = torch.arange(nseq + 1, dtype=torch.int32, device="cuda") * L
    q, k, v = qkv[:, :hq * d], qkv[:, hq * d:(hq + hkv) * d], qkv[:, (hq + hkv) * d:]
    ours = timed(lambda: cu.attn_prefill(q, k, v, out, cu_seq, nseq, L, hq, hkv, d, d ** -0.5))
    q4 = q.reshape(nseq, L, hq, d).transpose(1, 2)
    k4 = k.reshape(nseq, L, hkv, d).transpose(1, 2)
    v4 = v.reshape(nseq, L, hkv, d).transpose(1, 2)
    try:
        lib = timed(lambda: F.scaled_dot_product_attention(q4, k4, v4, is_causal=True, enable_gqa=(hq != hkv)))
    except Exception as e:  # noqa: BLE001
        lib = float("nan")
    flops = 4.0 * nseq * hq * d * L * L / 2  # causal: half of the S x S work, QK^T + PV
    print(f"nseq={nseq:3d} L={L:5d} hq={hq:2d} hkv={hkv:2d} d={d:3d} | ours {ours:9.1f} us = {flops / ours / 1e6:7.1f} TFLOP/s | "
          f"torch SDPA {lib:9.1f} us = {flops / lib / 1e6:7.1f} TFLOP/s | {lib / ours:.2f}x", flush=True)


if __name__ == "__main__":
    bench(16, 1024, 16, 16, 256)
    bench(16, 1024, 2, 2, 256)
    bench(16, 1024, 32, 32, 128)
    bench(4, 4096, 32, 32, 128)
    bench(2, 8192, 32, 32, 128)
    bench(16, 1024, 32, 32, 64)
    bench(16, 1024, 48, 1, 128)
