"""Headline benchmark: batched decode tokens/s on the GPT-J-6B architecture at TP = N.

    python bench.py                       # N=1, defaults that finish in a few minutes
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...  # the UNMODIFIED reference (baseline/_ref), same metric

Metric (BASELINE.json): decode tokens/s, device-timed (CUDA events), max over ranks; plus p50
TTFT / TPOT.  Config: GPT-J-6B shape (the largest architecture the reference can load, BASELINE.md
R1), batch 64, 1024-token synthetic prompts, random-init weights in HF safetensors layout, bf16
(ours) / fp16 (reference hard-codes fp16, generate.py:53).  A "step" is one decode step of the
whole batch = `batch` new tokens.  TP serving splits a fixed batch over N GPUs: strong scaling.

Timing rules: W >= 3 warm-up steps; every step streams > 12 GB / N of weights plus the KV cache
from HBM, far beyond the 126 MB L2 ("inputs larger than L2"); CUDA events on the launching stream,
barrier + synchronize on both sides; nvidia-smi clocks sampled during the timed region.
The `e2e` number goes through the public Engine.step() API: each step copies its inputs from pinned
host memory to the device and reads the sampled tokens back.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
    os.environ["NCCL_DEBUG"] = "WARN"  # keep stdout to the one JSON line (NCCL prints its version banner there)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--model", default="gptj-6b")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--prompt-len", type=int, default=1024)
    ap.add_argument("--dtype", default=None, help="bf16 | fp16 (default: bf16 ours, fp16 reference)")
    ap.add_argument("--ckpt-dir", default=os.environ.get("LLMSS_BENCH_CKPT", "/tmp/llmss_b200_bench"))
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--block-size", type=int, default=None, help="KV page size (default: engine's choice, 64 on CUDA)")
    ap.add_argument("--profile-ops", action="store_true", help="eager decode with per-op CUDA-event timing (diagnostic)")
    return ap.parse_args()


# ----------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region.

    Primary source is NVML (the library behind nvidia-smi) polled every 2 ms from a thread, so that even a
    ~100 ms timed region (TP=8) gets dozens of samples; `nvidia-smi -lms 100` runs alongside as the
    recipe's reference reading and is used alone if NVML cannot be loaded."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NVML_REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown"}

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self.proc = gpu_index, [], None
        self.nv_sm, self.nv_reasons, self.nv_max, self.nv_power = [], set(), None, []
        self._stop = threading.Event()
        self.nv_thread = None

    def _nvml_handle(self):
        import pynvml

        pynvml.nvmlInit()
        try:
            import torch

            pr = torch.cuda.get_device_properties(self.idx)
            bus = "%08X:%02X:%02X.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
            return pynvml, pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        except Exception:
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.idx)

    def _poll(self, nv, h):
        while not self._stop.is_set():
            try:
                self.nv_sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
                mask = int(get(h))
                for bit, name in self.NVML_REASONS.items():
                    if mask & bit:
                        self.nv_reasons.add(name)
                self.nv_power.append(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        try:
            nv, h = self._nvml_handle()
            self.nv_max = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self.nv_thread = threading.Thread(target=self._poll, args=(nv, h), daemon=True)
            self.nv_thread.start()
        except Exception:
            self.nv_thread = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        self._stop.set()
        if self.nv_thread is not None:
            self.nv_thread.join(timeout=1.0)
        if self.proc is None and not self.nv_sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        if self.proc is not None:
            if not self.nv_sm:
                time.sleep(0.15)
            self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) > 2 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        reasons = set(self.nv_reasons)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for n, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        allsm = self.nv_sm + sm
        if self.nv_max:
            mx.append(self.nv_max)
        out = {"sm_mhz": statistics.median(allsm) if allsm else None, "sm_max_mhz": max(mx) if mx else None,
               "reasons": sorted(reasons), "samples": len(allsm), "nvidia_smi_samples": len(sm)}
        if self.nv_power:
            out["power_w"] = round(statistics.median(self.nv_power), 1)
        return out


# ----------------------------------------------------------------------------- helpers
def dist_setup(n_gpus: int):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    assert world == n_gpus, f"--gpus {n_gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {n_gpus}"
    torch.cuda.set_device(local)
    return rank, world, local


def max_over_ranks(x: float, world: int) -> float:
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def fence(world: int):
    if world > 1:
        dist.barrier(device_ids=[torch.cuda.current_device()])
    torch.cuda.synchronize()


def ensure_checkpoint(args, rank: int, world: int) -> str:
    from llmss_b200.utils.checkpoint import write_random_checkpoint

    path = Path(args.ckpt_dir) / args.model
    if rank == 0:
        t0 = time.time()
        write_random_checkpoint(path, args.model, dtype=torch.float16, device="cuda")
        print(f"[bench] checkpoint {path} ready in {time.time() - t0:.1f}s", file=sys.stderr, flush=True)
    if world > 1:
        dist.barrier(device_ids=[torch.cuda.current_device()])
    return str(path)


def synthetic_prompts(batch: int, prompt_len: int, vocab: int):
    g = torch.Generator().manual_seed(1234)
    return torch.randint(10, vocab - 10, (batch, prompt_len), generator=g).tolist()


BASELINE_VALUE = None  # the reference publishes no number


def bench_config(args, world: int) -> dict:
    """The workload description -- byte-identical in both arms (implementation details go under "details")."""
    return {"model": f"{args.model} (random-init, HF safetensors layout)", "global_batch": args.batch,
            "seq_len": args.prompt_len, "prompt_len": args.prompt_len, "parallelism": f"tp{world}",
            "l2": "inputs larger than L2 (weights + KV streamed each step)"}


# ----------------------------------------------------------------------------- our arm
def run_ours(args, rank, world, local):
    from llmss_b200.engine.engine import Engine
    from llmss_b200.engine.runner import SamplingParams
    from llmss_b200.ops import _lib
    from llmss_b200.utils.dist import initialize_torch_distributed

    pg, _, _ = initialize_torch_distributed(timeout_s=600)
    path = ensure_checkpoint(args, rank, world)
    dtype = {"bf16": torch.bfloat16, "fp16": torch.float16}[args.dtype or "bf16"]
    K, Wm, B, S = args.steps, args.warmup, args.batch, args.prompt_len
    total_new = 1 + 2 * (K + Wm) + 8
    t0 = time.time()
    eng = Engine.from_pretrained(path, dtype=dtype, device=torch.device("cuda", local), process_group=pg,
                                 max_batch=B, max_seq_len=S + total_new + 16, use_cuda_graph=not args.no_graph,
                                 max_prefill_tokens=8192, symmetric_bytes=128 << 20, lamport_bytes=64 << 20,
                                 block_size=args.block_size)
    load_s = time.time() - t0
    runner = eng.runner
    prompts = synthetic_prompts(B, S, eng.core.spec.vocab)
    sp = SamplingParams(max_new_tokens=total_new, is_greedy=True)
    for p in prompts:
        eng.add_request(p, sp)

    # ---- prefill (TTFT): continuous-batching admission in chunks of max_prefill_tokens
    fence(world)
    ttfts = []
    t_pf0 = time.perf_counter()
    while eng.waiting:
        eng.step()
        torch.cuda.synchronize()
        ttfts.append(time.perf_counter() - t_pf0)
    prefill_s = time.perf_counter() - t_pf0
    assert len(eng.running) == B

    # ---- e2e: public API, host round trip every step (first call also binds + captures the graph)
    for _ in range(Wm):
        eng.step()
    fence(world)
    h2d0, d2h0 = runner.stats["h2d_bytes"], runner.stats["d2h_bytes"]
    step_times = []
    t_e0 = time.perf_counter()
    for _ in range(K):
        ts = time.perf_counter()
        outs = eng.step()
        step_times.append(time.perf_counter() - ts)
    fence(world)
    e2e_s = max_over_ranks(time.perf_counter() - t_e0, world)
    assert len(outs) == B
    h2d = (runner.stats["h2d_bytes"] - h2d0) // K
    d2h = (runner.stats["d2h_bytes"] - d2h0) // K

    # ---- device-timed decode: enqueue-only steps, tokens fed back on the device
    for _ in range(Wm):
        runner.decode_step(sync=False, host_tokens=False)
    fence(world)
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = _lib.LAUNCHES["count"]
    replays0 = runner.stats["graph_replays"]
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(K):
        runner.decode_step(sync=False, host_tokens=False)
    ev1.record()
    fence(world)
    clocks = sampler.stop()
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1), world)
    runner.sync_lengths()
    replays = runner.stats["graph_replays"] - replays0
    per_graph = runner.launches_per_graph.get(runner._bucket(B), 0)
    gpu_launches = replays * per_graph + (_lib.LAUNCHES["count"] - launches0)

    value = B * K / (dev_ms / 1e3)
    e2e_value = B * K / e2e_s
    if args.profile_ops:
        from llmss_b200.utils import tracing

        tracing.enable(profile=True)
        tracing.reset()
        runner.use_graph = False
        for _ in range(4):
            runner.decode_step(sync=False, host_tokens=False)
        rep = tracing.report()
        if rank == 0:
            print("[bench] per-op device time over 4 eager decode steps:", file=sys.stderr)
            for k, v in rep.items():
                print(f"[bench]   {k:28s} calls={v['calls']:5d} avg={v['avg_us']:9.2f} us total={v['total_us'] / 4:10.1f} us/step",
                      file=sys.stderr)
        runner.sync_lengths()
    out = {
        "metric": "decode_tokens_per_s", "value": round(value, 1), "unit": "tokens/s", "n_gpus": world, "steps": K,
        "warmup": Wm, "ms_per_step": round(dev_ms / K, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "bf16" if dtype == torch.bfloat16 else "fp16", "data": "synthetic",
        "impl": "ours",
        "config": bench_config(args, world),
        "details": {"cuda_graph": not args.no_graph, "fused_parallel_block": eng.core.fuse_parallel,
                    "symmetric_heap": getattr(eng.comm.symm, "mode", None), "nvls_multicast": bool(getattr(eng.comm.symm, "has_multicast", False)),
                    "heads_per_rank": eng.core.hq, "padded_heads": eng.core.heads_padded - eng.core.spec.n_heads},
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 1), "unit": "tokens/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "ms_per_step": round(1e3 * e2e_s / K, 4)},
        "gpu_launches": int(gpu_launches),
        "latency": {"ttft_p50_ms": round(1e3 * statistics.median(ttfts), 2), "prefill_total_ms": round(1e3 * prefill_s, 2),
                    "tpot_p50_ms": round(1e3 * statistics.median(step_times), 4)},
        "load_s": round(load_s, 1),
        "attn_split_tuning": runner.tuning_log.get(runner._bucket(B)),
    }
    if rank == 0:
        print(json.dumps(out), flush=True)


# ----------------------------------------------------------------------------- reference arm
def run_reference(args, rank, world, local):
    """The unmodified reference package (baseline/_ref/llmss) driven by a loop equivalent to its
    generate.py:102-145 (model -> logits -> rank-0 argmax -> broadcast), same checkpoint / shapes."""
    try:
        from baseline.ref_runner import build_reference_model, init_reference_dist

        init_reference_dist()  # the reference's own NCCL bring-up (dist.py:40-77)
    except Exception as e:  # noqa: BLE001
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": f"reference import failed: {e}"[:300]}), flush=True)
        return
    path = ensure_checkpoint(args, rank, world)
    K, Wm, B, S = args.steps, args.warmup, args.batch, args.prompt_len
    try:
        from transformers import AutoConfig

        mt = AutoConfig.from_pretrained(path).model_type
        if mt in ("gptj", "gpt_bigcode"):
            tier = "R1: unmodified reference model code"
            model, process_group, config = build_reference_model(path, local)
        else:
            # the reference has no model file for this family (custom_modeling/__init__.py:4-7): its own TP layer
            # library assembled to the same shapes (baseline/ref_r2.py)
            from baseline.ref_r2 import build_r2_model

            tier = "R2: reference library path at equal shapes (reference TP layers + ATen, no reference model file exists)"
            model, process_group, config = build_r2_model(path, local)
    except Exception as e:  # noqa: BLE001
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": f"reference model build failed: {e}"[:300]}), flush=True)
        return
    dev = torch.device("cuda", local)
    prompts = torch.tensor(synthetic_prompts(B, S, config.vocab_size), dtype=torch.int64, device=dev)

    def step(ids, past):
        with torch.no_grad():
            out = model(ids, past_key_values=past, use_cache=True)
        nxt = torch.zeros((B, 1), dtype=torch.int64, device=dev)  # generate.py:107
        if rank == 0:
            nxt = out.logits[:, -1, :].argmax(-1).unsqueeze(-1)     # generate.py:111
        if world > 1:
            dist.broadcast(nxt, src=0)                              # generate.py:144
        return nxt, out.past_key_values

    fence(world)
    t0 = time.perf_counter()
    ids, past = step(prompts, None)
    torch.cuda.synchronize()
    ttft = time.perf_counter() - t0
    for _ in range(Wm):
        ids, past = step(ids, past)
    # ---- e2e: the loop as a user of generate.py runs it, tokens read back on the host every step
    fence(world)
    step_times = []
    t_e0 = time.perf_counter()
    for _ in range(K):
        ts = time.perf_counter()
        ids, past = step(ids, past)
        _ = ids.cpu()
        step_times.append(time.perf_counter() - ts)
    fence(world)
    e2e_s = max_over_ranks(time.perf_counter() - t_e0, world)
    # ---- device-timed: no host read inside the event window
    fence(world)
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(K):
        ids, past = step(ids, past)
    ev1.record()
    fence(world)
    clocks = sampler.stop()
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1), world)
    value = B * K / (dev_ms / 1e3)
    out = {
        "metric": "decode_tokens_per_s", "value": round(value, 1), "unit": "tokens/s", "n_gpus": world, "steps": K,
        "warmup": Wm, "ms_per_step": round(dev_ms / K, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "fp16", "data": "synthetic", "impl": "reference",
        "config": bench_config(args, world),
        "details": {"tier": tier, "shims": "accelerate.init_empty_weights stub; PreTrainedModel.get_head_mask; eager attention"},
        "clocks": clocks,
        "e2e": {"value": round(B * K / e2e_s, 1), "unit": "tokens/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": B * 8, "ms_per_step": round(1e3 * e2e_s / K, 4)},
        "gpu_launches": 0,
        "latency": {"ttft_p50_ms": round(1e3 * ttft, 2), "tpot_p50_ms": round(1e3 * statistics.median(step_times), 4)},
    }
    if rank == 0:
        print(json.dumps(out), flush=True)


def main():
    args = parse()
    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device"}))
        return
    rank, world, local = dist_setup(args.gpus)
    if args.impl == "reference":
        run_reference(args, rank, world, local)
    else:
        run_ours(args, rank, world, local)
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
