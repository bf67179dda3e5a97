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
    symm = comm.attach_symmetric(128 << 20)
    rows = []
    shapes = [("gptj out+down", 4096, 20480), ("llama7b o_proj", 4096, 4096), ("llama7b down", 4096, 11008),
              ("llama13b down", 5120, 13824)]
    for name, N, Kfull in shapes:
        K = (Kfull // world) // 64 * 64
        for M in (1, 16, 64, 256):
            nvar = max(2, min(16, int(300e6 // (N * K * 2)) + 1))
            ws = [torch.randn(N, K, device=dev, dtype=torch.bfloat16) / math.sqrt(K * world) for _ in range(nvar)]
            x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
            y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
            plan = cu.gemm_plan(M, N, K, world)
            if plan.ar_bytes > symm.lam_bytes:
                continue
            i = [0]

            def fused():
                i[0] += 1
                cu.linear(x, ws[i[0] % nvar], symm=symm, out=y)

            def gemm_only():
                i[0] += 1
                cu.linear(x, ws[i[0] % nvar], out=y)

            def ours_unfused():
                i[0] += 1
                cu.linear(x, ws[i[0] % nvar], out=y)
                symm.all_reduce_(y)

            def ar_only():
                symm.all_reduce_(y)

            def reference_path():  # layers.py:175-179: F.linear (cuBLAS) then NCCL all_reduce
                i[0] += 1
                torch.mm(x, ws[i[0] % nvar].t(), out=y)
                dist.all_reduce(y)

            def nccl_only():
                dist.all_reduce(y)

            r = dict(name=name, M=M, N=N, K=K, tp=world, split=plan.split, fused_us=graph_time(fused),
                     gemm_us=graph_time(gemm_only), unfused_us=graph_time(ours_unfused), ar_us=graph_time(ar_only),
                     ref_us=graph_time(reference_path), nccl_us=graph_time(nccl_only))
            # roofline: slower of weight streaming at measured HBM peak and the all-reduce bytes over NVLink
            t_hbm = N * K * 2 / 6545.6e9 * 1e6
            t_link = (world - 1) * M * N * 2 / 770e9 * 1e6
            r["roofline_us"] = max(t_hbm, t_link)
            r["frac"] = r["roofline_us"] / r["fused_us"]
            rows.append(r)
            if rank == 0:
                print(f"{name:15s} tp={world} M={M:3d} N={N} K={K:5d} split={plan.split} | fused {r['fused_us']:7.1f} us | "
                      f"gemm {r['gemm_us']:6.1f} + one-shot AR {r['ar_us']:6.1f} = unfused {r['unfused_us']:7.1f} | "
                      f"cuBLAS+NCCL {r['ref_us']:7.1f} (NCCL alone {r['nccl_us']:6.1f}) | roofline {r['roofline_us']:5.1f} us "
                      f"-> {100 * r['frac']:4.1f}% | {r['ref_us'] / r['fused_us']:.2f}x vs reference path", flush=True)
    if rank == 0 and "--json" in sys.argv:
        json.dump(rows, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
