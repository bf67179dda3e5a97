"""Tensor parallelism without a cluster: real 2- and 4-rank gloo process groups on CPU.

Checks: TP=t logits == TP=1 logits for every family, the engine produces
identical tokens on every rank without a per-token broadcast, and the collective count per forward
is what the design promises (1 all-reduce per layer for the fused GPT-J block, 2 otherwise, + 1
all-gather for the head) -- vs. the reference's 1+2L / 2+2L all-reduces.
"""
import os
import socket
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

sys.path.insert(0, str(Path(__file__).parent))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, path, ids, out_file, fuse):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    os.environ.pop("DEBUG", None)
    from transformers import AutoConfig

    from llmss_b200.engine.engine import Engine
    from llmss_b200.engine.runner import SamplingParams
    from llmss_b200.models import MODEL_REGISTRY
    from llmss_b200.parallel.comm import comm_for
    from llmss_b200.utils.dist import initialize_torch_distributed
    from llmss_b200.utils.hub import weight_files
    from llmss_b200.utils.weights import Weights

    pg, r, w = initialize_torch_distributed()
    assert (r, w) == (rank, world) and dist.is_initialized()
    cfg = AutoConfig.from_pretrained(path)
    weights = Weights(weight_files(path), "cpu", torch.float32, pg)
    model = MODEL_REGISTRY[cfg.model_type](cfg, weights, fuse_parallel=fuse).eval()
    comm = comm_for(pg, torch.device("cpu"))
    comm.stats.update({k: 0 for k in comm.stats})
    out = model(ids, use_cache=True)
    stats = dict(comm.stats)
    step = model(ids[:, :1], past_key_values=out.past_key_values, use_cache=True)
    eng = Engine(model, max_batch=2, max_seq_len=48)
    toks = eng.generate([ids[0].tolist(), ids[1, :5].tolist()], SamplingParams(max_new_tokens=5, top_k=8, top_p=0.9, seed=3))
    torch.save({"logits": out.logits, "step": step.logits, "stats": stats, "toks": toks}, f"{out_file}.{rank}")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("family,world,fuse", [("gptj", 2, True), ("gptj", 2, False), ("gpt_bigcode", 2, True),
                                               ("gpt2", 2, True), ("opt", 4, True), ("llama", 2, True),
                                               ("llama", 4, True)])
def test_tp_equals_single(family, world, fuse, tmp_path):
    from tiny import make_tiny

    from llmss_b200.engine.engine import Engine
    from llmss_b200.engine.runner import SamplingParams
    from llmss_b200.models import MODEL_REGISTRY
    from llmss_b200.utils.dist import FakeGroup
    from llmss_b200.utils.hub import weight_files
    from llmss_b200.utils.weights import Weights
    from transformers import AutoConfig

    cfg, _ = make_tiny(family, tmp_path)
    path = str(tmp_path)
    conf = AutoConfig.from_pretrained(path)
    single = MODEL_REGISTRY[conf.model_type](conf, Weights(weight_files(path), "cpu", torch.float32, FakeGroup(0, 1))).eval()
    torch.manual_seed(5)
    ids = torch.randint(0, cfg.vocab_size, (2, 8))
    ref = single(ids, use_cache=True)
    ref_step = single(ids[:, :1], past_key_values=ref.past_key_values, use_cache=True)
    ref_toks = Engine(single, max_batch=2, max_seq_len=48).generate(
        [ids[0].tolist(), ids[1, :5].tolist()], SamplingParams(max_new_tokens=5, top_k=8, top_p=0.9, seed=3))

    out_file = str(tmp_path / "out")
    mp.spawn(_worker, args=(world, _free_port(), path, ids, out_file, fuse), nprocs=world, join=True)
    L = conf.n_layer if hasattr(conf, "n_layer") else conf.num_hidden_layers
    for r in range(world):
        got = torch.load(f"{out_file}.{r}")
        assert torch.allclose(got["logits"], ref.logits, atol=3e-4, rtol=1e-4), (family, r)
        assert torch.allclose(got["step"], ref_step.logits, atol=3e-4, rtol=1e-4), (family, r)
        assert got["toks"] == ref_toks, (family, r, got["toks"], ref_toks)  # every rank samples identically
        per_layer = 1 if (family == "gptj" and fuse) else 2
        assert got["stats"]["all_reduce"] == per_layer * L, got["stats"]
        assert got["stats"]["all_gather"] == 1 and got["stats"]["broadcast"] == 0
