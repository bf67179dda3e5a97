"""Engine on CPU (oracle ops, no CUDA graph): continuous batching must reproduce per-request greedy decoding."""
import torch

from llmss_b200.engine.engine import Engine
from llmss_b200.engine.runner import SamplingParams
from tiny import make_tiny


def _hf_greedy(hf, prompt, n):
    ids = torch.tensor([prompt])
    out = []
    with torch.no_grad():
        for _ in range(n):
            t = int(hf(ids).logits[0, -1].argmax())
            out.append(t)
            ids = torch.cat([ids, torch.tensor([[t]])], 1)
    return out


def test_engine_greedy_matches_hf(tmp_path):
    cfg, hf = make_tiny("gptj", tmp_path)
    eng = Engine.from_pretrained(str(tmp_path), dtype=torch.float32, device="cpu", max_batch=4, max_seq_len=64)
    torch.manual_seed(0)
    prompts = [torch.randint(0, cfg.vocab_size, (n,)).tolist() for n in (3, 7, 5, 9, 4, 6)]  # > max_batch: queueing
    sp = [SamplingParams(max_new_tokens=m, is_greedy=True) for m in (5, 3, 8, 4, 6, 2)]
    got = eng.generate(prompts, sp)
    for p, s, g in zip(prompts, sp, got):
        assert g == _hf_greedy(hf, p, s.max_new_tokens)
    assert eng.runner.alloc.num_free == eng.runner.alloc.num_blocks - 1  # every block returned


def test_engine_sampling_is_reproducible_and_respects_top_k(tmp_path):
    cfg, hf = make_tiny("llama", tmp_path)
    eng = Engine.from_pretrained(str(tmp_path), dtype=torch.float32, device="cpu", max_batch=2, max_seq_len=64)
    p = [[1, 2, 3, 4]]
    a = eng.generate(p, SamplingParams(max_new_tokens=6, top_k=1, top_p=1.0, temperature=0.7, seed=7))
    assert a[0] == _hf_greedy(hf, p[0], 6)  # top_k = 1 is greedy
    b = eng.generate(p, SamplingParams(max_new_tokens=6, top_k=20, top_p=0.9, seed=11))
    c = eng.generate(p, SamplingParams(max_new_tokens=6, top_k=20, top_p=0.9, seed=11))
    assert b == c


def test_max_seq_len_stops_generation(tmp_path):
    cfg, _ = make_tiny("gpt2", tmp_path)
    eng = Engine.from_pretrained(str(tmp_path), dtype=torch.float32, device="cpu", max_batch=2, max_seq_len=16)
    out = eng.generate([[5] * 10], SamplingParams(max_new_tokens=50, is_greedy=True))
    assert len(out[0]) == 6  # 10 prompt + 6 generated = 16 = max_seq_len (no sliding window)
