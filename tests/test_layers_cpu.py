"""Shard rules + TP layer API on CPU, single process (FakeGroup per rank)."""
import pytest
import torch
from safetensors.torch import save_file

from llmss_b200.utils.dist import FakeGroup, initialize_torch_distributed
from llmss_b200.utils.hub import weight_files
from llmss_b200.utils.layers import (FastLayerNorm, TensorParallelColumnLinear, TensorParallelEmbedding,
                                     TensorParallelHead, TensorParallelRowLinear)
from llmss_b200.utils.weights import Weights


@pytest.fixture
def ckpt(tmp_path):
    torch.manual_seed(0)
    t = {"a.weight": torch.randn(12, 8), "a.bias": torch.randn(12), "b.weight": torch.randn(12, 8),
         "r.weight": torch.randn(6, 16), "r.bias": torch.randn(6), "emb.weight": torch.randn(11, 4),
         "head.weight": torch.randn(10, 8), "ln.weight": torch.randn(8), "ln.bias": torch.randn(8),
         "q.idx": torch.arange(6, dtype=torch.int32)}
    save_file({k: v for k, v in list(t.items())[:5]}, str(tmp_path / "m-1.safetensors"))
    save_file({k: v for k, v in list(t.items())[5:]}, str(tmp_path / "m-2.safetensors"))
    return tmp_path, t


def W(path, rank, size, dtype=torch.float32):
    return Weights(weight_files(str(path)), "cpu", dtype, FakeGroup(rank, size))


def test_routing_and_dup_key_error(ckpt, tmp_path):
    path, t = ckpt
    w = W(path, 0, 1)
    assert set(w.routing) == set(t)
    assert w.get_shape("r.weight") == [6, 16]
    assert torch.equal(w.get_tensor("q.idx"), t["q.idx"])  # int payloads are not cast
    save_file({"a.weight": torch.zeros(1)}, str(path / "m-3.safetensors"))
    with pytest.raises(RuntimeError, match="multiple files"):
        W(path, 0, 1)


def test_aliases_and_missing(ckpt):
    path, t = ckpt
    w = Weights(weight_files(str(path)), "cpu", torch.float32, FakeGroup(0, 1), aliases={"x.weight": ["a.weight"]})
    assert torch.equal(w.get_tensor("x.weight"), t["a.weight"])
    with pytest.raises(RuntimeError, match="does not exist"):
        w.get_tensor("nope")


@pytest.mark.parametrize("size", [2, 3])
def test_column_row_rules(ckpt, size):
    path, t = ckpt
    if 12 % size:
        return
    cols, rows = [], []
    for r in range(size):
        w = W(path, r, size)
        c = TensorParallelColumnLinear.load(None, "a", w, bias=True)
        blk = 12 // size
        assert torch.equal(c.linear.weight, t["a.weight"][r * blk:(r + 1) * blk])
        assert torch.equal(c.linear.bias, t["a.bias"][r * blk:(r + 1) * blk])
        cols.append(c)
        if 16 % size == 0:
            rl = TensorParallelRowLinear.load(None, "r", w, bias=True)
            kb = 16 // size
            assert torch.equal(rl.linear.weight, t["r.weight"][:, r * kb:(r + 1) * kb])
            assert (rl.linear.bias is not None) == (r == 0)  # bias on rank 0 only
            rows.append(rl)
    multi = TensorParallelColumnLinear.load_multi(None, ["a", "b"], W(path, 1, size), bias=False, dim=0)
    blk = 12 // size
    assert torch.equal(multi.linear.weight, torch.cat([t["a.weight"][blk:2 * blk], t["b.weight"][blk:2 * blk]]))
    # sum of row-parallel partials == full linear
    if rows:
        x = torch.randn(3, 16)
        kb = 16 // size
        total = sum(F_lin(rows[r], x[:, r * kb:(r + 1) * kb]) for r in range(size))
        assert torch.allclose(total, x @ t["r.weight"].t() + t["r.bias"], atol=1e-5)


def F_lin(layer, x):
    return torch.nn.functional.linear(x, layer.linear.weight, layer.linear.bias)


def test_indivisible_shard_asserts(ckpt):
    path, _ = ckpt
    with pytest.raises(AssertionError):
        W(path, 0, 5).get_sharded("a.weight", 0)


def test_embedding_floor_and_null_row(ckpt):
    path, t = ckpt
    embs = [TensorParallelEmbedding("emb", W(path, r, 2), reduce=False) for r in range(2)]
    assert embs[0].weight.shape == (6, 4) and embs[0].null_idx == 5  # 11 // 2 = 5 rows + 1 zero row
    ids = torch.tensor([[0, 4, 5, 9, 10]])
    total = embs[0](ids) + embs[1](ids)
    want = t["emb.weight"][ids]
    assert torch.allclose(total[0, :4], want[0, :4])
    assert torch.equal(total[0, 4], torch.zeros(4))  # reference quirk Q10: id 10 falls off the floor-divided shards
    # the engine-side rule (ceil) keeps every row
    lo, hi, blk = W(path, 1, 2).shard_bounds(11, ceil=True)
    assert (lo, hi, blk) == (6, 11, 6)


def test_head_shard_or_replicate(ckpt):
    path, t = ckpt
    h = TensorParallelHead.load(None, "head", W(path, 1, 2))
    assert h.should_gather and torch.equal(h.linear.weight, t["head.weight"][5:])
    h3 = TensorParallelHead.load(None, "head", W(path, 1, 3))  # 10 % 3 != 0 -> replicated, no gather
    assert not h3.should_gather and h3.linear.weight.shape == (10, 8)


def test_layernorm_loaders_and_fast_ln(ckpt):
    path, t = ckpt
    w = W(path, 0, 1)
    ln = torch.nn.LayerNorm.load("ln", w, 1e-5)
    x = torch.randn(5, 8)
    assert torch.allclose(ln(x), torch.nn.functional.layer_norm(x, (8,), t["ln.weight"], t["ln.bias"], 1e-5))
    assert torch.nn.LayerNorm.load_no_bias("ln", w, 1e-5).bias is None
    fln = FastLayerNorm.load("ln", w, 1e-5)
    res = torch.randn(5, 8)
    y, r = fln(x, res)
    assert torch.allclose(r, x + res) and torch.allclose(y, ln(x + res), atol=1e-5)


def test_fake_group_and_init(monkeypatch):
    monkeypatch.setenv("WORLD_SIZE", "1")
    monkeypatch.setenv("RANK", "0")
    g, rank, ws = initialize_torch_distributed()
    assert isinstance(g, FakeGroup) and (rank, ws) == (0, 1)
    monkeypatch.setenv("WORLD_SIZE", "4")
    monkeypatch.setenv("RANK", "2")
    monkeypatch.setenv("DEBUG", "1")
    g, rank, ws = initialize_torch_distributed()
    assert isinstance(g, FakeGroup) and g.rank() == 2 and g.size() == 4
    g.allreduce([torch.ones(1)]).wait()
    g.barrier().wait()


def test_weight_files_local_and_missing(ckpt, tmp_path):
    path, _ = ckpt
    assert [p.name for p in weight_files(str(path))] == ["m-1.safetensors", "m-2.safetensors"]
    empty = tmp_path / "empty"
    empty.mkdir()
    with pytest.raises(FileNotFoundError):
        weight_files(str(empty))


def test_fused_allreduce_mode_selection():
    """Which exchange a row-parallel GEMM fuses into its epilogue (ops._fused_ar_mode): pure host logic, pinned here
    with a stand-in for the GEMM planner.  Thresholds are the measured ones (profiles/bench_gemm_ar.md)."""
    from types import SimpleNamespace

    from llmss_b200 import ops

    def plan_for(split, bn=64):
        def gemm_plan(M, N, K, nranks):
            tiles = -(-N // 128) * -(-M // bn)
            return SimpleNamespace(split=split, tiles=tiles, bn=bn, ar_bytes=nranks * tiles * 128 * bn * 2)
        return SimpleNamespace(gemm_plan=gemm_plan)

    def symm(t, lam=64 << 20):
        return SimpleNamespace(nranks=t, lam_bytes=lam)

    cluster, streamk = plan_for(4), plan_for(0)
    # stream-K shapes never fuse (measured slower fused than unfused)
    assert ops._fused_ar_mode(streamk, symm(8), 64, 4096, 2560, 2) == 0
    # TP=8, GPT-J out+down, M=64: 512 KB message -> two-shot Lamport exchange; M=16 (128 KB) -> one-shot push
    assert ops._fused_ar_mode(cluster, symm(8), 64, 4096, 2560, 2) == 6
    assert ops._fused_ar_mode(cluster, symm(8), 16, 4096, 2560, 2) == 5
    # two ranks: the one-shot push up to 1 MB, two-shot from there
    assert ops._fused_ar_mode(cluster, symm(2), 64, 4096, 10240, 2) == 5
    assert ops._fused_ar_mode(cluster, symm(2), 128, 4096, 10240, 2) == 6
    # a Lamport area too small for either form -> plain GEMM + standalone collective
    assert ops._fused_ar_mode(cluster, symm(8, lam=1 << 20), 64, 4096, 2560, 2) == 0
    # prefill-sized messages (above LLMSS_FUSED_TWOSHOT_MAX_BYTES) are left to the standalone all-reduce
    assert ops._fused_ar_mode(plan_for(4, bn=256), symm(8), 8192, 4096, 2560, 2) == 0
