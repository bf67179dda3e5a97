"""CPU oracle: every family vs. HF transformers on tiny random checkpoints."""
import pytest
import torch

from llmss_b200.models import MODEL_REGISTRY
from llmss_b200.utils.dist import FakeGroup
from llmss_b200.utils.hub import weight_files
from llmss_b200.utils.weights import Weights
from tiny import make_tiny

FAMILIES = ["gptj", "gpt_bigcode", "gpt2", "opt", "llama"]


def _load(path, fuse=True, dtype=torch.float32):
    from transformers import AutoConfig

    cfg = AutoConfig.from_pretrained(str(path))
    w = Weights(weight_files(str(path)), device="cpu", dtype=dtype, process_group=FakeGroup(0, 1))
    return MODEL_REGISTRY[cfg.model_type](cfg, w, fuse_parallel=fuse).eval()


@pytest.mark.parametrize("family", FAMILIES)
def test_matches_hf_prefill_and_cached_decode(family, tmp_path):
    cfg, hf = make_tiny(family, tmp_path)
    ours = _load(tmp_path)
    torch.manual_seed(1)
    ids = torch.randint(0, cfg.vocab_size, (2, 9))
    with torch.no_grad():
        ref = hf(ids).logits
    out = ours(ids, use_cache=True)
    assert out.logits.shape == ref.shape
    assert torch.allclose(out.logits, ref, atol=2e-4, rtol=1e-4), (out.logits - ref).abs().max()
    # cached single-token steps == full recompute
    full = torch.cat([ids, torch.randint(0, cfg.vocab_size, (2, 3))], dim=1)
    past = out.past_key_values
    for t in range(9, 12):
        step = ours(full[:, t:t + 1], past_key_values=past, use_cache=True)
        past = step.past_key_values
        with torch.no_grad():
            want = hf(full[:, :t + 1]).logits[:, -1]
        assert torch.allclose(step.logits[:, -1], want, atol=2e-4, rtol=1e-4), (family, t)


def test_gptj_unfused_equals_fused(tmp_path):
    cfg, _ = make_tiny("gptj", tmp_path)
    a, b = _load(tmp_path, fuse=True), _load(tmp_path, fuse=False)
    ids = torch.randint(0, cfg.vocab_size, (1, 7))
    assert torch.allclose(a(ids).logits, b(ids).logits, atol=1e-5)


def test_multi_token_query_with_context(tmp_path):
    """Chunked prefill (q_len > 1 with prior context) goes through the generic attention path."""
    cfg, hf = make_tiny("llama", tmp_path)
    ours = _load(tmp_path)
    ids = torch.randint(0, cfg.vocab_size, (2, 10))
    o1 = ours(ids[:, :6], use_cache=True)
    o2 = ours(ids[:, 6:], past_key_values=o1.past_key_values, use_cache=True)
    with torch.no_grad():
        ref = hf(ids).logits
    assert torch.allclose(torch.cat([o1.logits, o2.logits], 1), ref, atol=2e-4, rtol=1e-4)
