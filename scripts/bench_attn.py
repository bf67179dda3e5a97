"""Device-timed decode-attention micro-benchmark (graph of back-to-back launches, KV >> L2)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llmss_b200.ops import cuda_impl as cu

try:
    PEAK_BW = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] * 1e9
except Exception:
    PEAK_BW = 6545.6e9


def bench(B, hq, hkv, d, ctx, splits=None, dtype=torch.bfloat16, reps=6, bs=16):
    maxb = (ctx + bs - 1) // bs
    nb = 1 + B * maxb
    kc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=dtype)
    vc = torch.randn(nb, hkv, bs, d, device="cuda", dtype=dtype)
    bt = (torch.arange(B * maxb, device="cuda", dtype=torch.int32) + 1).view(B, maxb)
    sl = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
    q = torch.randn(B, hq * d, device="cuda", dtype=dtype)
    out = torch.empty_like(q)
    sp = cu.decode_splits(B, hq, hkv, d, ctx, dtype, bs) if splits is None else splits
    fn = lambda: cu.attn_decode(q, kc, vc, bt, sl, out, hq, hkv, d, d ** -0.5, sp)  # noqa: E731
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3 / reps)
    t = sorted(ts)[len(ts) // 2]
    bytes_ = 2.0 * B * ctx * hkv * d * 2
    print(f"B={B:3d} hq={hq:2d} hkv={hkv:2d} d={d:3d} ctx={ctx:5d} bs={bs:2d} splits={sp:2d} | {t:8.1f} us | "
          f"{bytes_ / t / 1e3:7.1f} GB/s = {100 * bytes_ / (t * 1e-6) / PEAK_BW:5.1f}% of measured HBM peak", flush=True)
    return t


if __name__ == "__main__":
    for sp in (None, 1, 2, 3, 5, 8):
        bench(64, 16, 16, 256, 1040, sp)
    bench(64, 8, 8, 256, 1040)
    bench(64, 4, 4, 256, 1040)
    bench(64, 2, 2, 256, 1040)
    bench(64, 32, 32, 128, 1040)
    bench(64, 4, 4, 128, 1040)
    bench(1, 16, 16, 256, 2000)
    bench(8, 16, 16, 256, 2000)
    bench(64, 48, 1, 128, 1040)
    bench(64, 6, 1, 128, 1040)
    for sp in (1, 2, 4, 8):
        bench(64, 48, 1, 128, 1040, sp)
    for bs in (16, 64):
        bench(64, 48, 1, 128, 1040, bs=bs)
        bench(64, 32, 8, 128, 1040, bs=bs)   # llama-3-style GQA
        bench(64, 32, 8, 128, 4096, bs=bs)
        bench(64, 8, 2, 128, 4096, bs=bs)
    if os.environ.get("LLMSS_BENCH_TC_MHA"):
        cu._TC_MIN_GROUP = 1
        for bs in (16, 64):
            bench(64, 16, 16, 256, 1040, bs=bs)
            bench(64, 32, 32, 128, 1040, bs=bs)
    print("balanced (splits=0) vs picked, 64-token pages, TP=1/2/4/8 head counts of GPT-J and a GQA model")
    for hq_, hkv_, d_, ctx_ in ((16, 16, 256, 1040), (8, 8, 256, 1040), (4, 4, 256, 1040), (2, 2, 256, 1040), (32, 8, 128, 4096),
                                (4, 1, 128, 4096)):
        bench(64, hq_, hkv_, d_, ctx_, bs=64)
        bench(64, hq_, hkv_, d_, ctx_, 0, bs=64)
    print("head_dim 64 (OPT / GPT-2 family), 64-token pages")
    bench(64, 32, 32, 64, 1040, bs=64)
    bench(64, 16, 16, 64, 1040, bs=64)
    bench(64, 25, 25, 64, 1040, bs=64)
