"""The shard-summary sampler (what a vocab-parallel CUDA sampler exchanges: local max, local sum-exp, local
top-k candidates with ties) must pick exactly the tokens of the full-logits sampler."""
import pytest
import torch

from llmss_b200.ops import reference as R


def _case(B, V, t, seed, ties=False):
    g = torch.Generator().manual_seed(seed)
    Vb = -(-V // t)
    logits = torch.randn(B, Vb * t, generator=g) * 3
    if ties:
        logits = torch.round(logits * 2) / 2  # many exact ties, inside and across shards
    logits[:, V:] = 7.0  # padding columns of the last shard must never be chosen
    return logits, [logits[:, r * Vb:(r + 1) * Vb].clone() for r in range(t)]


@pytest.mark.parametrize("t", [1, 2, 4, 8])
@pytest.mark.parametrize("ties", [False, True])
def test_shard_summaries_reproduce_full_sampler(t, ties):
    B, V = 24, 1003  # V not divisible by t: ceil-div shards with padding
    logits, shards = _case(B, V, t, seed=10 + t, ties=ties)
    greedy = torch.tensor([1, 0, 0, 0] * (B // 4), dtype=torch.int32)
    temp = torch.tensor([1.0, 0.7, 1.0, 0.5, 0.9, 1.0] * (B // 6))
    top_k = torch.tensor([50, 50, 0, 5, 1, 64, 17, 3] * (B // 8), dtype=torch.int32)
    top_p = torch.tensor([0.95, 0.95, 1.0, 0.5, 0.9, 1.0, 0.3, 1.0] * (B // 8))
    for step in range(5):
        rng_a = torch.stack([torch.arange(B) * 7 + 1, torch.full((B,), step)], dim=1)
        rng_b = rng_a.clone()
        want = R.sample(logits.clone(), greedy, temp, top_k, top_p, rng_a, vocab=V)
        got = R.sample_vocab_parallel(shards, greedy, temp, top_k, top_p, rng_b, vocab=V, k_summary=64)
        assert torch.equal(rng_a, rng_b)
        assert got.tolist() == want.tolist(), (t, ties, step)
        assert int(got.max()) < V


def test_rows_beyond_the_summary_size_are_refused():
    logits, shards = _case(4, 4000, 2, seed=3)
    z = torch.zeros(4, dtype=torch.int32)
    with pytest.raises(ValueError):
        # top_k = 0 with top_p < 1 searches the 1024 most probable tokens: not coverable by 64 candidates per rank
        R.sample_vocab_parallel(shards, z, torch.ones(4), z, torch.full((4,), 0.9),
                                torch.zeros(4, 2, dtype=torch.int64), vocab=4000, k_summary=64)
