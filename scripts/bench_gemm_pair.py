"""CTA-pair GEMM kernel (tcgen05 cta_group::2, csrc/gemm_tcgen05.cu: gemm_pair_kernel): correctness against an fp32
reference for every epilogue flavour, then device-timed A/B against the single-CTA kernels and cuBLAS.

    python scripts/bench_gemm_pair.py [--check-only] [--json out.json]

Timing as in scripts/bench_gemm.py: weights cycled through a pool larger than L2, launches captured in one CUDA
graph, CUDA events around replays, median of 10.
"""
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from llmss_b200.ops import _lib
from llmss_b200.ops import cuda_impl as cu
from llmss_b200.ops import reference as R
from bench_gemm import PEAK_BW, PEAK_FLOPS, timeit  # noqa: E402

lib = _lib.load()


def set_pair(mode: int):
    lib.llmss_gemm_set_pair(mode)


def check(M, N, K, split, dtype=torch.bfloat16, flavour="plain"):
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N + K + split)
    x = (torch.randn(M, K, device="cuda", generator=g) * 0.5).to(dtype)
    w = (torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K)).to(dtype)
    bias = torch.randn(N, device="cuda", generator=g).to(dtype)
    kw, ref = {}, x.float() @ w.float().t()
    if flavour == "bias_gelu_from":       # GPT-J fused QKV + fc_in: activation only for features >= act_start
        start = (N // 256) * 128
        kw = dict(bias=bias, act=R.ACT_GELU_TANH, act_start=start)
        ref = ref + bias.float()
        ref[:, start:] = F.gelu(ref[:, start:], approximate="tanh")
    elif flavour == "gated":              # SwiGLU on interleaved (gate, up) rows
        kw = dict(act=R.ACT_SILU, gated=True)
        ref = F.silu(ref[:, 0::2]) * ref[:, 1::2]
    elif flavour == "residual":
        res = torch.randn(M, N, device="cuda", generator=g).to(dtype)
        kw = dict(bias=bias, residual=res)
        ref = ref + bias.float() + res.float()
    elif flavour == "f32":
        kw = dict(out_dtype=torch.float32)
    set_pair(split)
    got = cu.linear(x, w, **kw)
    set_pair(0)
    base = cu.linear(x, w, **kw)
    torch.cuda.synchronize()
    err = (got.float() - ref).abs().max().item()
    err_base = (base.float() - ref).abs().max().item()
    scale = ref.abs().max().item()
    same = (got.float() - base.float()).abs().max().item()
    ok = err <= max(2.5 * err_base, 0.01 * scale)
    print(f"check M={M:4d} N={N:6d} K={K:6d} pairs/cluster={split} {str(dtype)[6:]:8s} {flavour:14s} max|err| {err:.4f} "
          f"(single-CTA kernel {err_base:.4f}, scale {scale:.1f}, pair vs single {same:.4f}) {'OK' if ok else 'FAIL'}", flush=True)
    return ok


def bench(M, N, K, dtype=torch.bfloat16):
    wbytes = N * K * 2
    nvar = max(2, min(64, int(400e6 // wbytes) + 1))
    ws = [torch.randn(N, K, device="cuda", dtype=dtype) / math.sqrt(K) for _ in range(nvar)]
    x = torch.randn(M, K, device="cuda", dtype=dtype)
    out = torch.empty(M, N, device="cuda", dtype=dtype)
    row = dict(M=M, N=N, K=K)
    set_pair(0)
    row["single_us"] = timeit(lambda i: cu.linear(x, ws[i], out=out), nvar)[1]
    for s in (1, 2, 4):
        if s > 1 and ((-(-N // 256)) * (-(-M // 256)) * 2 * s > 148 or K // 64 < 2 * s):
            continue
        set_pair(s)
        row[f"pair{s}_us"] = timeit(lambda i: cu.linear(x, ws[i], out=out), nvar)[1]
    set_pair(0)
    bo = torch.empty(M, N, device="cuda", dtype=dtype)
    row["cublas_us"] = timeit(lambda i: torch.mm(x, ws[i].t(), out=bo), nvar)[1]
    roof = max((wbytes + M * K * 2 + M * N * 2) / PEAK_BW, 2.0 * M * N * K / PEAK_FLOPS) * 1e6
    row["roofline_us"] = roof
    best = min(v for k, v in row.items() if k.startswith("pair"))
    row["best_pair_us"] = best
    print(f"M={M:5d} N={N:6d} K={K:6d} | single-CTA {row['single_us']:7.1f} us | pairs " +
          " ".join(f"S={s}: {row[f'pair{s}_us']:7.1f}" for s in (1, 2, 4) if f"pair{s}_us" in row) +
          f" | cuBLAS {row['cublas_us']:7.1f} us | roofline {roof:6.1f} us | best pair {best / row['single_us']:.2f}x of single, "
          f"{row['cublas_us'] / best:.2f}x cuBLAS", flush=True)
    return row


if __name__ == "__main__":
    ok = True
    if "--threshold" in sys.argv:
        check = lambda *a, **k: True  # noqa: E731 - correctness of this build is covered by tests/test_gpu_kernels.py
    for split in (1, 2, 4):
        ok &= check(256, 512, 512, split)
    for flavour in ("bias_gelu_from", "gated", "residual", "f32"):
        ok &= check(256, 1024, 1024, 2, flavour=flavour)
    ok &= check(200, 4096 + 128, 4096 + 64, 4)                 # ragged rows, odd feature-tile count, K tail
    ok &= check(129, 3584, 4096, 2, dtype=torch.float16)
    ok &= check(700, 768, 2048, 1, flavour="residual")         # several token tiles
    print("PAIR_CHECK", "OK" if ok else "FAILED", flush=True)
    if "--check-only" in sys.argv or not ok:
        sys.exit(0 if ok else 1)
    rows = []
    if "--threshold" in sys.argv:   # where does the persistent pair schedule start to pay? (pair tiles: 16 .. 48)
        for M, N, K in ((512, 2048, 4096), (256, 6144, 4096), (512, 3584, 4096), (256, 8192, 4096), (512, 4096, 4096),
                        (256, 10240, 4096), (384, 5120, 5120), (768, 4096, 4096), (256, 8192, 8192), (512, 4096, 11008)):
            rows.append(bench(M, N, K))
        json.dump(rows, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1) if "--json" in sys.argv else None
        sys.exit(0)
    for M in ((256,) if "--large" in sys.argv else (256, 192)):
        for N, K in (((12288, 4096), (28672, 4096)) if "--large" in sys.argv else
                     ((4096, 4096), (12288, 4096), (16384, 4096), (4096, 16384), (28672, 4096), (3584, 4096), (4096, 2560),
                      (50432, 4096), (4096, 20480))):
            rows.append(bench(M, N, K))
    for M, N, K in ((512, 12288, 4096), (1024, 4096, 4096), (2048, 12288, 4096), (2048, 4096, 16384), (8192, 12288, 4096),
                    (8192, 4096, 4096), (8192, 16384, 4096), (8192, 4096, 16384), (8192, 28672, 4096)):
        rows.append(bench(M, N, K))
    if "--json" in sys.argv:
        json.dump(rows, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)
