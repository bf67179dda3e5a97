"""Graph-timed fused GEMM -> all-reduce vs its parts vs the reference path (cuBLAS + NCCL), max over ranks.

    torchrun --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_gemm_ar.py
Shapes: the row-parallel GEMMs of GPT-J-6B (fused out_proj+fc_out: N=4096, K=20480/N) and Llama-2 7B/13B
(o_proj, down_proj) for decode batches M=1..256 (BASELINE config 5).
"""
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from llmss_b200 import ops
from llmss_b200.ops import cuda_impl as cu
from llmss_b200.parallel.comm import comm_for
from llmss_b200.utils.dist import initialize_torch_distributed


def graph_time(fn, reps=20, iters=5):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    ts = []
    for _ in range(iters):
        torch.cuda.synchronize()
        dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) * 1e3 / reps], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ts.append(t.item())
    return sorted(ts)[len(ts) // 2]


def main():
    pg, rank, world = initialize_torch_distributed(timeout_s=300)
    dev = torch.device("cuda", torch.cuda.current_device())
    comm = comm_for(pg, dev)
    symm = comm.attach_symmetric(256 << 20, lamport_bytes=64 << 20)
    if rank == 0:
        print(f"symmetric heap: mode={symm.mode} multicast={symm.has_multicast} log={symm.bringup_log}", flush=True)
    rows = []
    shapes = [("gptj out+down", 4096, 20480), ("llama7b o_proj", 4096, 4096), ("llama7b down", 4096, 11008),
              ("llama13b down", 5120, 13824)]
    Ms = (1, 16, 64, 128, 256) + ((8192,) if "--prefill" in sys.argv else ())
    if "--quick" in sys.argv:
        shapes, Ms = shapes[:1] + shapes[3:], (16, 64, 256)
    nan = float("nan")
    for name, N, Kfull in shapes:
        K = (Kfull // world) // 64 * 64
        for M in Ms:
            nvar = max(2, min(16, int(300e6 // (N * K * 2)) + 1))
            ws = [torch.randn(N, K, device=dev, dtype=torch.bfloat16) / math.sqrt(K * world) for _ in range(nvar)]
            x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
            y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
            plan = cu.gemm_plan(M, N, K, world)
            i = [0]

            def w():
                i[0] += 1
                return ws[i[0] % nvar]

            msg = M * N * 2
            can_push = plan.ar_bytes <= symm.lam_bytes
            can_nvls = symm.has_multicast and plan.split > 0 and plan.tiles * 128 * plan.bn * 2 <= min(symm.lam_bytes, symm.buf_bytes)
            r = dict(name=name, M=M, N=N, K=K, tp=world, split=plan.split, bn=plan.bn, tiles=plan.tiles)
            r["gemm_us"] = graph_time(lambda: cu.linear(x, w(), out=y))
            r["fused_push_us"] = graph_time(lambda: cu.linear(x, w(), symm=symm, out=y, collective=5)) if can_push else nan
            r["fused_nvls_us"] = graph_time(lambda: cu.linear(x, w(), symm=symm, out=y, collective=4)) if can_nvls else nan
            can_two = plan.split > 0 and (world + 1) * plan.tiles * 128 * plan.bn * 2 <= symm.lam_bytes
            r["fused_two_us"] = graph_time(lambda: cu.linear(x, w(), symm=symm, out=y, collective=6)) if can_two else nan
            r["ar_twolam_us"] = graph_time(lambda: symm.all_reduce_(y, algo=4)) if (world + 1) * msg <= symm.lam_bytes else nan
            r["ar_oneshot_us"] = graph_time(lambda: symm.all_reduce_(y, algo=1)) if msg * world <= symm.lam_bytes and msg <= (1 << 20) else nan
            r["ar_twoshot_us"] = graph_time(lambda: symm.all_reduce_(y, algo=2)) if (msg // 16) % world == 0 and 2 * msg <= symm.buf_bytes else nan
            r["ar_nvls_us"] = graph_time(lambda: symm.all_reduce_(y, algo=3)) if symm.has_multicast and 2 * msg <= symm.buf_bytes else nan

            def product_path():  # what ops.linear(..., comm=comm) does for this shape
                ops.linear(x, w(), comm=comm)

            r["ours_us"] = graph_time(product_path)
            r["mode"] = ops._fused_ar_mode(cu, symm, M, N, K, 2)

            def reference_path():  # layers.py:175-179: F.linear (cuBLAS) then NCCL all_reduce
                torch.mm(x, w().t(), out=y)
                dist.all_reduce(y)

            r["ref_us"] = graph_time(reference_path)
            r["nccl_us"] = graph_time(lambda: dist.all_reduce(y))
            r["cublas_us"] = graph_time(lambda: torch.mm(x, w().t(), out=y))
            # roofline: slowest of weight streaming at the measured HBM peak, the flops at the measured bf16 peak and
            # the bytes that must cross NVLink (in-switch reduction: 2 (t-1)/t x message per direction) at 770 GB/s
            t_hbm = (N * K * 2 + M * K * 2 + M * N * 2) / 6545.6e9 * 1e6
            t_flop = 2.0 * M * N * K / 1654.1e12 * 1e6
            t_link = 2.0 * (world - 1) / world * msg / 770e9 * 1e6
            r["roofline_us"] = max(t_hbm, t_flop, t_link)
            r["frac"] = r["roofline_us"] / r["ours_us"]
            rows.append(r)
            if rank == 0:
                print(f"{name:15s} tp={world} M={M:4d} N={N} K={K:5d} split={plan.split} | gemm {r['gemm_us']:7.1f} | fused push {r['fused_push_us']:7.1f} "
                      f"two {r['fused_two_us']:7.1f} nvls {r['fused_nvls_us']:7.1f} | AR 1shot {r['ar_oneshot_us']:7.1f} 2lam {r['ar_twolam_us']:7.1f} 2flag {r['ar_twoshot_us']:7.1f} nvls {r['ar_nvls_us']:7.1f} | "
                      f"ours(mode {r['mode']}) {r['ours_us']:7.1f} | cuBLAS {r['cublas_us']:7.1f} + NCCL {r['nccl_us']:6.1f} = {r['ref_us']:7.1f} | "
                      f"roofline {r['roofline_us']:6.1f} -> {100 * r['frac']:4.1f}% | {r['ref_us'] / r['ours_us']:.2f}x vs ref", flush=True)
    if rank == 0 and "--json" in sys.argv:
        json.dump(rows, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
