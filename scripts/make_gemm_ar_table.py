"""Render profiles/raw/gemm_ar_tp*.json (scripts/bench_gemm_ar.py --json) into profiles/bench_gemm_ar.md."""
import glob
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def f(x):
    return "-" if x is None or x != x else f"{x:.1f}"


def main():
    out = ["# Row-parallel GEMM -> all-reduce (BASELINE.json config 5): every row, losing ones included", "",
           "`scripts/bench_gemm_ar.py` under `torchrun`, one rank per B200: each variant is captured as a CUDA graph of 20",
           "back-to-back launches (weights cycled through a pool larger than L2), timed with CUDA events, max over ranks,",
           "median of 5 replays.  All times in microseconds.  Raw rows: `raw/gemm_ar_tp*.json`.", "",
           "* **gemm**: the tcgen05 GEMM alone (no collective).  **fused push / two / nvls**: the same GEMM with the",
           "  all-reduce in its epilogue -- one-shot Lamport push (t-1 x the message on the wire, one hop), two-shot",
           "  Lamport exchange (reduce-scatter by peer stores + `multimem.st` broadcast, ~1 x the message, two hops), and",
           "  the flag-synchronised in-switch reduction (`multimem.ld_reduce`).",
           "* **AR ...**: the standalone collectives on the same message (1shot / 2lam / 2flag / nvls as above).",
           "* **ours**: what `ops.linear(x, w, comm=...)` actually runs for the shape (`mode`: 5 = fused push, 6 = fused",
           "  two-shot, 0 = GEMM + standalone collective).  **ref**: the reference's path, `F.linear` (cuBLAS) followed",
           "  by `torch.distributed.all_reduce` (NCCL), `layers.py:175-179`.",
           "* **roofline**: the slowest of (weights + activations) / 6545.6 GB/s measured HBM, flops / 1654.1 TFLOP/s measured",
           "  bf16, and 2 (t-1)/t x message / 770 GB/s measured NVLink per direction; **%** = roofline / ours.", ""]
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "raw", "gemm_ar_tp*.json"))):
        rows = json.load(open(path))
        if not rows:
            continue
        out += [f"## TP = {rows[0]['tp']}", "",
                "| shape | M | N | K/rank | split | gemm | fused push | fused two | fused nvls | AR 1shot | AR 2lam | AR 2flag | AR nvls | **ours** (mode) | cuBLAS | NCCL | **ref** | roofline | % | ours vs ref |",
                "|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
        for r in rows:
            out.append(f"| {r['name']} | {r['M']} | {r['N']} | {r['K']} | {r['split']} | {f(r['gemm_us'])} | {f(r['fused_push_us'])} | "
                       f"{f(r.get('fused_two_us'))} | {f(r['fused_nvls_us'])} | {f(r['ar_oneshot_us'])} | {f(r.get('ar_twolam_us'))} | "
                       f"{f(r['ar_twoshot_us'])} | {f(r['ar_nvls_us'])} | **{f(r['ours_us'])}** ({r['mode']}) | {f(r['cublas_us'])} | "
                       f"{f(r['nccl_us'])} | **{f(r['ref_us'])}** | {f(r['roofline_us'])} | {100 * r['frac']:.0f} | {r['ref_us'] / r['ours_us']:.2f}x |")
        worst = min(rows, key=lambda r: r["ref_us"] / r["ours_us"])
        out += ["", f"Slowest row relative to the reference path: {worst['name']} M={worst['M']}: {worst['ref_us'] / worst['ours_us']:.2f}x.", ""]
    open(os.path.join(ROOT, "profiles", "bench_gemm_ar.md"), "w").write("\n".join(out) + "\n")
    print("wrote profiles/bench_gemm_ar.md")


if __name__ == "__main__":
    main()
