"""Multi-GPU correctness of the symmetric-memory kernels (run under torchrun, one rank per GPU).

    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 scripts/mgpu_check.py

Checks, against NCCL / single-GPU results:
  1. one-shot and two-shot all-reduce (bf16 / fp16 / fp32, many sizes, back-to-back calls);
  2. all-gather;
  3. fused tcgen05 GEMM -> all-reduce (bias + residual epilogue), full tiles and stream-K partial tiles;
  4. the same kernels replayed from a CUDA graph (epoch / parity protocol);
  5. end to end: TP=N greedy tokens of the engine == TP=1 tokens (rank-replicated sampling).
"""
import math
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from llmss_b200.ops import cuda_impl as cu
from llmss_b200.parallel.comm import comm_for
from llmss_b200.utils.dist import initialize_torch_distributed


def log(*a):
    if dist.get_rank() == 0:
        print(*a, flush=True)


def main():
    pg, rank, world = initialize_torch_distributed(timeout_s=120)
    dev = torch.device("cuda", torch.cuda.current_device())
    comm = comm_for(pg, dev)
    symm = comm.attach_symmetric(64 << 20, lamport_bytes=max(16, 8 * world) << 20)  # bulk + Lamport area (grows with the rank count)
    assert symm is not None
    ok = True
    only_engine = "--only-engine" in sys.argv
    log(f"symmetric heap: mode={symm.mode} multicast={symm.has_multicast} log={symm.bringup_log}")
    algos = (1, 2, 3, 4) if symm.has_multicast else (1, 2, 4)

    if not only_engine:
        # ---- 1. all-reduce
        for dtype in (torch.bfloat16, torch.float16, torch.float32):
            for n in (8, 1024, 4096 * 64, 4096 * 64 + 8, 1 << 20, 3 << 20):
                for algo in algos:
                    nbytes = n * torch.empty((), dtype=dtype).element_size()
                    if algo == 2 and (nbytes // 16) % world != 0:
                        continue
                    if (algo == 1 and nbytes * world > symm.lam_bytes) or (algo in (2, 3) and nbytes * 2 > symm.buf_bytes):
                        continue
                    if algo == 4 and nbytes * (world + 1) > symm.lam_bytes:
                        continue
                    torch.manual_seed(1000 + rank)
                    x = torch.randn(n, device=dev, dtype=dtype)
                    want = x.clone()
                    dist.all_reduce(want)
                    for rep in range(3):  # back-to-back: exercises epoch parity / buffer reuse
                        y = x.clone()
                        symm.all_reduce_(y, algo=algo)
                        torch.cuda.synchronize()
                        tol = 2e-2 if dtype != torch.float32 else 1e-5
                        good = torch.allclose(y.float(), want.float(), atol=tol * math.sqrt(world), rtol=tol)
                        if not good:
                            ok = False
                            log(f"FAIL allreduce dtype={dtype} n={n} algo={algo} rep={rep} max err {(y.float() - want.float()).abs().max().item()}")
                    # every rank must hold identical bits
                    gathered = [torch.empty_like(y) for _ in range(world)]
                    dist.all_gather(gathered, y)
                    if not all(torch.equal(gathered[0], g) for g in gathered):
                        ok = False
                        log(f"FAIL allreduce not bit-identical across ranks dtype={dtype} n={n} algo={algo}")
        log("allreduce checks done", "OK" if ok else "FAILED")

        # ---- 2. all-gather
        x = torch.full((1000, 8), float(rank), device=dev, dtype=torch.float32)
        g = symm.all_gather(x)
        torch.cuda.synchronize()
        for r in range(world):
            if not torch.all(g[r] == r):
                ok = False
                log("FAIL allgather slot", r)
        # column gather (vocab-parallel logits): cat(all_gather(x), dim=-1) in one pass, odd row counts / sizes
        for rows, cols in ((64, 6304), (5, 36), (1, 4), (33, 25216)):
            xc = (torch.arange(rows * cols, device=dev, dtype=torch.float32).view(rows, cols) + 1e6 * rank)
            gc = symm.all_gather_cols(xc)
            torch.cuda.synchronize()
            for r in range(world):
                want = torch.arange(rows * cols, device=dev, dtype=torch.float32).view(rows, cols) + 1e6 * r
                if not torch.equal(gc[:, r * cols:(r + 1) * cols], want):
                    ok = False
                    log("FAIL allgather_cols", rows, cols, "slot", r)
        for _ in range(20):  # back-to-back calls: parity double buffering
            gc = symm.all_gather_cols(xc)
        torch.cuda.synchronize()
        if not torch.equal(gc[:, rank * cols:(rank + 1) * cols], xc):
            ok = False
            log("FAIL allgather_cols repeat")
        log("allgather done", "OK" if ok else "FAILED")

        # ---- 3. fused GEMM -> all-reduce
        for (M, N, K) in [(1, 512, 256), (16, 4096, 2560), (64, 4096, 2560), (64, 4096, 20480 // world), (40, 1000, 328), (256, 2048, 512)]:
            for dtype in (torch.bfloat16, torch.float16):
                torch.manual_seed(7 + rank)
                xr = torch.randn(M, K, device=dev, dtype=dtype)
                wr = torch.randn(N, K, device=dev, dtype=dtype) / math.sqrt(K * world)
                torch.manual_seed(99)
                bias = torch.randn(N, device=dev, dtype=dtype)
                res = torch.randn(M, N, device=dev, dtype=dtype)
                part = (xr.float() @ wr.float().t())
                want = part.clone()
                dist.all_reduce(want)
                want = want + bias.float() + res.float()
                plan = cu.gemm_plan(M, N, K, world)
                if plan.ar_bytes > symm.lam_bytes:
                    continue
                modes = [5] + ([4] if (symm.has_multicast and plan.split > 0) else []) + ([6] if plan.split > 0 else [])
                for mode in modes:
                    for rep in range(3):
                        y = cu.linear(xr, wr, bias, residual=res, symm=symm, collective=mode)
                        torch.cuda.synchronize()
                        err = (y.float() - want).abs().max().item()
                        if err > 0.08:
                            ok = False
                            log(f"FAIL gemm+AR M={M} N={N} K={K} {dtype} mode={mode} rep={rep} max err {err}")
                    gathered = [torch.empty_like(y) for _ in range(world)]
                    dist.all_gather(gathered, y)
                    if not all(torch.equal(gathered[0], gg) for gg in gathered):
                        ok = False
                        log(f"FAIL gemm+AR not bit-identical across ranks M={M} N={N} K={K} mode={mode}")
        log("fused GEMM+AR done", "OK" if ok else "FAILED")

        # ---- 3b. fused GEMM -> reduce-scatter (K2) and all-gather -> GEMM (K3), prefill-size token counts
        for (M, N, K) in [(world * 256, 1024, 512), (world * 512, 4096, 1024), (world * 128, 768, 2048)]:
            torch.manual_seed(11 + rank)
            xr = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
            wr = torch.randn(N, K, device=dev, dtype=torch.bfloat16) / math.sqrt(K * world)
            torch.manual_seed(5)
            bias = torch.randn(N, device=dev, dtype=torch.bfloat16)
            res = torch.randn(M, N, device=dev, dtype=torch.bfloat16)
            full = xr.float() @ wr.float().t()
            dist.all_reduce(full)
            rows = M // world
            want = full[rank * rows:(rank + 1) * rows] + bias.float() + res[rank * rows:(rank + 1) * rows].float()
            for rep in range(2):
                got = cu.linear_reduce_scatter(xr, wr, symm, bias=bias, residual_shard=res[rank * rows:(rank + 1) * rows])
                torch.cuda.synchronize()
                err = (got.float() - want).abs().max().item()
                if err > 0.1:
                    ok = False
                    log(f"FAIL gemm+RS M={M} N={N} K={K} rep={rep} max err {err}")
            # all-gather -> GEMM: every rank contributes `rows` token rows, weights are rank-local (column parallel)
            torch.manual_seed(21 + rank)
            xs = torch.randn(rows, K, device=dev, dtype=torch.bfloat16)
            wc = torch.randn(N, K, device=dev, dtype=torch.bfloat16) / math.sqrt(K)
            bc = torch.randn(N, device=dev, dtype=torch.bfloat16)
            gathered = [torch.empty_like(xs) for _ in range(world)]
            dist.all_gather(gathered, xs)
            xfull = torch.cat(gathered, 0)
            from llmss_b200.ops import reference as R
            want2 = R.linear(xfull, wc, bc, act=R.ACT_GELU_TANH, out_dtype=torch.float32)
            for rep in range(2):
                got2 = cu.allgather_linear(xs, wc, symm, bias=bc, act=R.ACT_GELU_TANH)
                torch.cuda.synchronize()
                err = (got2.float() - want2).abs().max().item()
                if err > 0.08:
                    ok = False
                    log(f"FAIL AG+gemm M={M} N={N} K={K} rep={rep} max err {err}")
        log("fused GEMM+reduce-scatter / all-gather+GEMM done", "OK" if ok else "FAILED")

        # ---- 4. CUDA graph replay of a GEMM+AR / all-reduce chain
        M, N, K = 64, 4096, 2048
        torch.manual_seed(3 + rank)
        xr = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        wr = torch.randn(N, K, device=dev, dtype=torch.bfloat16) / math.sqrt(K * world)
        z = torch.randn(4096 * 8, device=dev, dtype=torch.bfloat16)
        want = xr.float() @ wr.float().t()
        dist.all_reduce(want)
        zw = z.clone()
        dist.all_reduce(zw)
        stream = torch.cuda.Stream()
        stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(stream):
            for _ in range(2):
                y = cu.linear(xr, wr, symm=symm)
                zz = z.clone()
                symm.all_reduce_(zz)
        torch.cuda.current_stream().wait_stream(stream)
        torch.cuda.synchronize()
        dist.barrier()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            for _ in range(5):
                y = cu.linear(xr, wr, symm=symm)
                zz = z.clone()
                symm.all_reduce_(zz)
        for rep in range(4):
            graph.replay()
        torch.cuda.synchronize()
        e1 = (y.float() - want).abs().max().item()
        e2 = (zz.float() - zw.float()).abs().max().item()
        if e1 > 0.08 or e2 > 0.1:
            ok = False
            log(f"FAIL graph replay: gemm err {e1}, allreduce err {e2}")
        log(f"graph replay done (epoch now {symm.epoch()})")

        # ---- 4b. protocol stress: randomised per-rank arrival skew, alternating kernels / sizes / parities.
        # (compute-sanitizer's racecheck cannot see inter-GPU races)
        n_stress = int(sys.argv[sys.argv.index("--stress") + 1]) if "--stress" in sys.argv else 60
        import random
        rnd = random.Random(1234 + rank)
        cases = []
        for n in (64, 4096, 65536 + 8):
            torch.manual_seed(500 + rank)
            x = torch.randn(n, device=dev, dtype=torch.bfloat16)
            w_ = x.clone()
            dist.all_reduce(w_)
            cases.append(("ar", x, w_))
        torch.manual_seed(600 + rank)
        xg = torch.randn(32, 1024, device=dev, dtype=torch.bfloat16)
        wg = torch.randn(4096, 1024, device=dev, dtype=torch.bfloat16) / math.sqrt(1024 * world)
        stress_split = cu.gemm_plan(32, 4096, 1024, world).split > 0
        stress_nvls = symm.has_multicast and stress_split
        eg = xg.float() @ wg.float().t()
        dist.all_reduce(eg)
        cases.append(("gemm", (xg, wg), eg))
        bad = 0
        for it in range(n_stress):
            kind, inp, want_ = cases[(it * 7 + 3) % len(cases)]
            if rnd.random() < 0.5:
                torch.cuda._sleep(int(rnd.random() * 400000))  # up to ~0.2 ms of skew on this rank only
            if kind == "ar":
                y_ = inp.clone()
                symm.all_reduce_(y_, algo=((3 if symm.has_multicast else 4) if it % 2 else (4 if it % 4 else 1)) if (it % 3) else 0)
                e_ = (y_.float() - want_.float()).abs().max().item()
                tol_ = 0.06 * math.sqrt(world)
            else:
                y_ = cu.linear(inp[0], inp[1], symm=symm, collective=(4 if stress_nvls else 6) if it % 2 else (6 if (stress_split and it % 4) else 5))
                e_ = (y_.float() - want_).abs().max().item()
                tol_ = 0.08
            if e_ > tol_:
                bad += 1
        torch.cuda.synchronize()
        if bad:
            ok = False
            log(f"FAIL stress: {bad}/{n_stress} iterations wrong")
        log(f"protocol stress ({n_stress} skewed iterations) done", "OK" if bad == 0 else "FAILED")

        # ---- latency numbers (device-timed, max over ranks) vs NCCL
        for nbytes in (8 << 10, 64 << 10, 512 << 10):
            x = torch.randn(nbytes // 2, device=dev, dtype=torch.bfloat16)

            def timed(fn, iters=50):
                for _ in range(5):
                    fn()
                torch.cuda.synchronize()
                dist.barrier()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(iters):
                    fn()
                b.record()
                torch.cuda.synchronize()
                t = torch.tensor([a.elapsed_time(b) * 1e3 / iters], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                return t.item()

            ours = timed(lambda: symm.all_reduce_(x, algo=1))
            nvls = timed(lambda: symm.all_reduce_(x, algo=3)) if symm.has_multicast else float("nan")
            two = timed(lambda: symm.all_reduce_(x, algo=4))
            nccl = timed(lambda: dist.all_reduce(x))
            log(f"all-reduce {nbytes >> 10:4d} KiB x{world}: one-shot P2P {ours:7.1f} us | two-shot Lamport {two:7.1f} us | NVLS {nvls:7.1f} us | NCCL {nccl:7.1f} us")

    # ---- 5. engine TP=N vs TP=1 for the three sharding regimes: MHA (gptj), GQA with sharded kv heads (llama),
    #         MQA with replicated kv (gpt_bigcode)
    from transformers import AutoConfig

    from llmss_b200.engine.engine import Engine
    from llmss_b200.engine.runner import SamplingParams
    from llmss_b200.models import MODEL_REGISTRY
    from llmss_b200.utils.checkpoint import write_random_checkpoint
    from llmss_b200.utils.dist import FakeGroup
    from llmss_b200.utils.hub import weight_files
    from llmss_b200.utils.weights import Weights

    g = torch.Generator().manual_seed(77)
    # short prompts (fused all-reduce path) + long ones (>= tp*64 packed tokens: sequence-parallel prefill path)
    prompts = [[3, 14, 15, 92, 65], [35, 89, 79], torch.randint(1, 500, (150,), generator=g).tolist(),
               torch.randint(1, 500, (97,), generator=g).tolist()]
    n_new = 16
    for preset in ("gptj-tiny8", "llama-tiny8", "bigcode-tiny8"):
        path = f"/tmp/llmss_mgpu_{preset}"
        if rank == 0:
            write_random_checkpoint(path, preset, dtype=torch.float32, device="cpu", std=0.05)
        dist.barrier()
        for sampling in (False, True, "plain"):
            sp = SamplingParams(max_new_tokens=n_new, is_greedy=not sampling, top_k=20, top_p=0.9, temperature=0.8, seed=5)
            if sampling == "plain":   # whole-distribution rows: owner pick + 16-byte gather (vocab-parallel mode 3)
                sp = SamplingParams(max_new_tokens=n_new, is_greedy=False, top_k=0, top_p=1.0, temperature=0.9, seed=5)
            tp = Engine.from_pretrained(path, dtype=torch.bfloat16, device=dev, process_group=pg, max_batch=4, max_seq_len=192)
            got = tp.generate(prompts[:2], sp)
            tp.core.sp_prefill = True  # long prompts: exercise the sequence-parallel prefill dataflow end to end
            got = got + tp.generate(prompts[2:], sp)
            bad_tok = [t for o_ in got for t in o_ if not (0 <= t < 512)]
            if bad_tok:
                ok = False
                log(f"FAIL {preset} sampling={sampling}: token ids out of range: {bad_tok[:8]} in {got}")
                got = [[min(max(t, 0), 511) for t in o_] for o_ in got]
            all_out = [None] * world
            dist.all_gather_object(all_out, got)
            if any(o != all_out[0] for o in all_out):
                ok = False
                log(f"FAIL {preset} sampling={sampling}: ranks disagree on the generated tokens")
            want_mode = {False: 1, True: 2, "plain": 3}[sampling]
            if not any(k[1] == want_mode for k in tp.runner.graphs):
                ok = False
                log(f"FAIL {preset} sampling={sampling}: vocab-parallel mode {want_mode} was not the one captured: {list(tp.runner.graphs)}")
            if sampling:
                continue   # sampled tokens are compared across ranks only (vocab-parallel sampler, same RNG stream)
            single = Engine.from_pretrained(path, dtype=torch.bfloat16, device=dev, process_group=FakeGroup(0, 1),
                                            max_batch=4, max_seq_len=192)
            want_t = single.generate(prompts[:2], sp) + single.generate(prompts[2:], sp)
            same = sum(a == b for a, b in zip(got, want_t))
            # bf16 near-ties may flip between summation orders: every TP token must be within rounding of the fp32
            # oracle's argmax given the same prefix
            if rank == 0:
                conf = AutoConfig.from_pretrained(path)
                oracle = MODEL_REGISTRY[conf.model_type](conf, Weights(weight_files(path), "cpu", torch.float32, FakeGroup(0, 1))).eval()
                worst = 0.0
                for p_, o_ in zip(prompts, got):
                    with torch.no_grad():
                        lg = oracle(torch.tensor([list(p_) + list(o_)])).logits[0].float()
                    for t_, tok in enumerate(o_):
                        row = lg[len(p_) + t_ - 1]
                        worst = max(worst, (row.max() - row[tok]).item() / max(1.0, row.abs().max().item()))
                if worst > 0.03:
                    ok = False
                    log(f"FAIL {preset}: a TP={world} token is {worst:.3f} x max|logit| below the oracle's argmax")
                log(f"engine {preset} TP={world}: {same}/{len(prompts)} prompts identical to TP=1 over {n_new} tokens, "
                    f"worst gap to the fp32 oracle's argmax {worst:.4f} x max|logit|; first: {got[0][:8]}")
            del single
        del tp
        torch.cuda.empty_cache()
    log("engine TP checks done | collectives:", comm.stats)

    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    log("MGPU_CHECK", "OK" if flag.item() == 0 else "FAILED")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 0 else 1)


if __name__ == "__main__":
    main()
