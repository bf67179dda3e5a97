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


def test_sampling_does_not_depend_on_the_batch_slot(tmp_path):
    """The sampler's RNG is keyed on (request seed, tokens sampled so far), not on the row of the running batch:
    a request draws the same tokens alone, behind other requests, or after its neighbours changed."""
    cfg, _ = make_tiny("llama", tmp_path)
    sp = SamplingParams(max_new_tokens=8, top_k=30, top_p=0.95, temperature=0.9, seed=1234)
    other = SamplingParams(max_new_tokens=3, top_k=10, top_p=0.9, seed=99)
    probe = [7, 8, 9, 10, 11]
    eng = Engine.from_pretrained(str(tmp_path), dtype=torch.float32, device="cpu", max_batch=4, max_seq_len=64)
    alone = eng.generate([probe], sp)[0]
    crowded = eng.generate([[1, 2, 3], [4, 5], probe, [6, 6, 6, 6]], [other, other, sp, other])
    assert crowded[2] == alone           # slot 2 of 4, neighbours finish early (slots are re-bound mid-way)
    first = eng.generate([probe, [3, 2, 1]], [sp, other])
    assert first[0] == alone             # slot 0
    # without an explicit seed the stream comes from the request id: two identical requests differ from each other
    nos = SamplingParams(max_new_tokens=12, top_k=0, top_p=1.0, temperature=1.0)
    a, b = eng.generate([probe, probe], [nos, nos])
    assert a != b


def test_admission_reserves_worst_case_and_rejects_the_impossible(tmp_path):
    """Blocks are promised at admission (prompt + max_new), so a running sequence can never fail to grow; a request
    that cannot fit even an empty cache is reported as finished with token -1 instead of blocking the queue."""
    cfg, _ = make_tiny("gpt2", tmp_path)
    # 5 usable blocks of 16 tokens: two requests of 20 + 20 tokens (3 blocks each) cannot run together
    eng = Engine.from_pretrained(str(tmp_path), dtype=torch.float32, device="cpu", max_batch=4, max_seq_len=64,
                                 block_size=16, num_blocks=6)
    sp = SamplingParams(max_new_tokens=20, is_greedy=True)
    a = eng.add_request([5] * 20, sp)
    b = eng.add_request([6] * 20, sp)
    outs = {}
    for o in eng.step():                         # prefill: only `a` is admitted, `b` waits for blocks
        outs.setdefault(o.request_id, []).append(o.token)
    assert [s.seq_id for s in eng.running] == [a] and [s.seq_id for s in eng.waiting] == [b]
    assert eng.runner._outstanding == 1          # 2 blocks allocated for the prompt, the third is promised
    too_big = eng.add_request([7] * 60, SamplingParams(max_new_tokens=60, is_greedy=True))  # 64 tokens = 4 blocks: fits
    huge = eng.add_request([7] * 10, SamplingParams(max_new_tokens=20, is_greedy=True))
    while eng.has_work():
        for o in eng.step():
            outs.setdefault(o.request_id, []).append(o.token)
    assert len(outs[a]) == 20 and len(outs[b]) == 20 and len(outs[huge]) == 20
    assert len(outs[too_big]) == 4               # stops at max_seq_len = 64
    assert eng.runner._outstanding == 0 and eng.runner.alloc.num_free == 5
    eng2 = Engine.from_pretrained(str(tmp_path), dtype=torch.float32, device="cpu", max_batch=4, max_seq_len=64,
                                  block_size=16, num_blocks=3)
    never = eng2.add_request([7] * 40, SamplingParams(max_new_tokens=8, is_greedy=True))  # 3 blocks > 2 usable
    ok = eng2.add_request([7] * 10, SamplingParams(max_new_tokens=4, is_greedy=True))
    got = {}
    while eng2.has_work():
        for o in eng2.step():
            got.setdefault(o.request_id, []).append(o.token)
    assert got[never] == [-1] and len(got[ok]) == 4


def test_chunked_prefill_matches_whole_prompt_prefill(tmp_path):
    """Prompts fed in rounds of 5 tokens (queries over prior context) generate what whole-prompt prefill generates."""
    cfg, hf = make_tiny("llama", tmp_path)
    torch.manual_seed(1)
    prompts = [torch.randint(0, cfg.vocab_size, (n,)).tolist() for n in (17, 3, 11, 5)]
    sp = SamplingParams(max_new_tokens=5, is_greedy=True)
    whole = Engine.from_pretrained(str(tmp_path), dtype=torch.float32, device="cpu", max_batch=4, max_seq_len=64)
    chunked = Engine.from_pretrained(str(tmp_path), dtype=torch.float32, device="cpu", max_batch=4, max_seq_len=64,
                                     prefill_chunk=5)
    a, b = whole.generate(prompts, sp), chunked.generate(prompts, sp)
    assert a == b
    assert a[0] == _hf_greedy(hf, prompts[0], 5)


def test_no_cache_mode_matches_cached_generation(tmp_path):
    """The reference's `--use_cache`-off loop (whole sequence recomputed every step) must give the cached tokens,
    greedy and sampled (same per-request RNG stream)."""
    cfg, hf = make_tiny("gptj", tmp_path)
    torch.manual_seed(2)
    prompts = [torch.randint(0, cfg.vocab_size, (n,)).tolist() for n in (6, 3, 9)]
    eng = Engine.from_pretrained(str(tmp_path), dtype=torch.float32, device="cpu", max_batch=4, max_seq_len=64)
    g = SamplingParams(max_new_tokens=6, is_greedy=True)
    assert eng.generate_no_cache(prompts, g) == eng.generate(prompts, g)
    assert eng.generate_no_cache(prompts, g)[0] == _hf_greedy(hf, prompts[0], 6)
    sp = SamplingParams(max_new_tokens=6, top_k=20, top_p=0.9, temperature=0.8, seed=21)
    assert eng.generate_no_cache(prompts, sp) == eng.generate(prompts, sp)
    assert eng.runner.alloc.num_free == eng.runner.alloc.num_blocks - 1


def test_sliding_window_matches_block_window_oracle(tmp_path):
    """Past max_seq_len the engine keeps generating over its most recent tokens: whole cache blocks are evicted from
    the front, positions stay absolute (reference quirk Q7, generate.py:133-144, done properly).  Oracle: HF GPT-J in
    ONE pass over the whole token stream with a 4-D mask in which token q sees keys [window_start(q), q] -- exactly what
    a cache that evicts blocks produces, including the hidden states of tokens computed before an eviction."""
    import sys
    from pathlib import Path

    sys.path.insert(0, str(Path(__file__).parent))
    from tiny import make_tiny
    from transformers import AutoConfig, AutoModelForCausalLM

    cfg, _ = make_tiny("gptj", tmp_path)
    hc = AutoConfig.from_pretrained(str(tmp_path))
    hc.n_positions = hc.max_position_embeddings = 512          # the oracle's rotary table must reach past the window
    hf = AutoModelForCausalLM.from_pretrained(str(tmp_path), config=hc, torch_dtype=torch.float32, attn_implementation="eager").eval()
    W, bs, n_new = 64, 16, 70
    eng = Engine.from_pretrained(str(tmp_path), device="cpu", max_batch=2, max_seq_len=W, block_size=bs, sliding_window=True)
    assert eng.runner.sliding
    g = torch.Generator().manual_seed(1)
    prompts = [torch.randint(1, cfg.vocab_size, (n,), generator=g).tolist() for n in (40, 9)]
    outs = eng.generate(prompts, SamplingParams(max_new_tokens=n_new, is_greedy=True))
    assert [len(o) for o in outs] == [n_new, n_new]                      # 110 and 79 tokens through a 64-token window
    assert eng.runner.alloc.num_free == eng.runner.alloc.num_blocks - 1 and eng.runner._outstanding == 0
    for p, o in zip(prompts, outs):
        toks = p + o
        T = len(toks) - 1
        q = torch.arange(T)
        start = torch.clamp(((q + 1 - W + bs - 1) // bs) * bs, min=0)    # first key still cached when token q was fed
        start[: len(p)] = 0                                              # the prompt is prefilled under plain causal attention
        k = torch.arange(T)
        seen = (k[None, :] <= q[:, None]) & (k[None, :] >= start[:, None])
        neg = torch.finfo(torch.float32).min
        with torch.no_grad():
            win = hf(input_ids=torch.tensor([toks[:-1]]), attention_mask=torch.zeros(T, T).masked_fill(~seen, neg)[None, None],
                     position_ids=q[None]).logits[0]
            full = hf(input_ids=torch.tensor([toks[:-1]]), position_ids=q[None]).logits[0]
        rows = win[len(p) - 1:]
        for i, t in enumerate(o):
            gap = (rows[i].max() - rows[i][t]).item()
            assert gap <= 1e-3 * max(1.0, rows[i].abs().max().item()), (len(p), i, t, int(rows[i].argmax()), gap)
        assert (win[W + 1:] - full[W + 1:]).abs().max().item() > 1e-3      # the window is not a no-op
    # without the switch a sequence stops at max_seq_len; learned-position families never slide
    stop = Engine.from_pretrained(str(tmp_path), device="cpu", max_batch=2, max_seq_len=W, block_size=bs)
    assert len(stop.generate([prompts[0]], SamplingParams(max_new_tokens=n_new, is_greedy=True))[0]) == W - len(prompts[0])
    cfg2, _ = make_tiny("gpt2", tmp_path / "gpt2")
    learned = Engine.from_pretrained(str(tmp_path / "gpt2"), device="cpu", max_batch=1, max_seq_len=32, block_size=bs, sliding_window=True)
    assert not learned.runner.sliding
    assert len(learned.generate([prompts[1]], SamplingParams(max_new_tokens=n_new, is_greedy=True))[0]) == 32 - len(prompts[1])


def test_prefix_cache_reuses_prompt_blocks(tmp_path):
    """Prompts that start with the same tokens share the cache blocks of that prefix (reference-counted, LRU-evicted);
    only the rest is computed -- queries over prior context -- and the tokens are those of an engine without the cache."""
    from llmss_b200.utils.checkpoint import write_random_checkpoint

    path = str(write_random_checkpoint(tmp_path, "gptj-tiny", dtype=torch.float32, device="cpu", std=0.08))
    g = torch.Generator().manual_seed(9)
    system = torch.randint(1, 500, (40,), generator=g).tolist()                     # 2.5 blocks of 16
    tails = [torch.randint(1, 500, (n,), generator=g).tolist() for n in (7, 1, 30, 12)]
    prompts = [system + t for t in tails] + [system[:32], system[:16] + tails[2]]   # incl. a prompt that IS a cached prefix
    sp = SamplingParams(max_new_tokens=6, is_greedy=True)
    plain = Engine.from_pretrained(path, device="cpu", max_batch=4, max_seq_len=128, block_size=16)
    want = [plain.generate([p], sp)[0] for p in prompts]
    eng = Engine.from_pretrained(path, device="cpu", max_batch=4, max_seq_len=128, block_size=16, prefix_cache=True)
    r = eng.runner
    got = [eng.generate([p], sp)[0] for p in prompts]                               # one after the other: later ones hit
    assert got == want
    # 2 blocks of the system prompt for requests 2-4, 1 block for system[:32] (its own last block is recomputed) and for the last
    assert r.stats["prefix_hit_tokens"] == 3 * 32 + 16 + 16
    assert r.stats["prefill_tokens"] == sum(len(p) for p in prompts) - r.stats["prefix_hit_tokens"]
    assert r.alloc.num_free == r.alloc.num_blocks - 1 and r._outstanding == 0 and len(r.alloc.evictable) > 0
    # all at once: requests admitted together compute the prefix once each (nothing to hit yet) but stay correct,
    # and a second wave hits what the first registered
    both = eng.generate(prompts[:4], sp) + eng.generate(prompts[:4], sp)
    assert both == want[:4] + want[:4]
    # chunked prefill on top of a cached prefix
    ch = Engine.from_pretrained(path, device="cpu", max_batch=4, max_seq_len=128, block_size=16, prefix_cache=True, prefill_chunk=8)
    assert [ch.generate([p], sp)[0] for p in prompts] == want
    # a cache too small to keep everything: least recently used prefixes are evicted, results stay right
    small = Engine.from_pretrained(path, device="cpu", max_batch=2, max_seq_len=128, block_size=16, prefix_cache=True, num_blocks=8)
    other = [torch.randint(1, 500, (50,), generator=g).tolist() for _ in range(3)]
    seq = [prompts[0], other[0], other[1], prompts[2], other[2], prompts[0]]
    assert [small.generate([p], sp)[0] for p in seq] == [plain.generate([p], sp)[0] for p in seq]
    a = small.runner.alloc
    assert a.num_free == a.num_blocks - 1 and all(x == 0 for x in a.ref) and len(small.runner._nodes) == len(a.kept)


def test_block_allocator_refcounts_and_lru():
    from llmss_b200.engine.kv_cache import BlockAllocator

    evicted = []
    a = BlockAllocator(6)                       # blocks 1..5 usable, 0 is the padding target
    a.on_evict = evicted.append
    x = a.alloc(3)
    assert 0 not in x and a.num_free == 2 and [a.ref[b] for b in x] == [1, 1, 1]
    a.keep(x[0]); a.keep(x[1])                  # registered in a prefix index
    a.share(x[0])                               # a second sequence holds it
    a.release(x)                                # first holder leaves: x[0] still held, x[1] evictable, x[2] free
    assert a.ref[x[0]] == 1 and x[1] in a.evictable and x[2] in a.free and a.num_free == 4
    a.release([x[0]])
    assert list(a.evictable) == [x[1], x[0]] and a.num_free == 5          # least recently used first
    a.share(x[1])                               # a hit revives an evictable block
    assert x[1] not in a.evictable and a.num_free == 4
    y = a.alloc(4)                              # drains the free list, then evicts x[0]
    assert evicted == [x[0]] and x[0] in y and x[0] not in a.kept and a.num_free == 0
    try:
        a.alloc(1)
        raise AssertionError("expected MemoryError")
    except MemoryError:
        pass
    a.release(y); a.release([x[1]])
    assert a.num_free == 5 and all(r == 0 for r in a.ref)


def test_prefix_cache_covers_generated_tokens_for_follow_up_turns(tmp_path):
    """A follow-up prompt that repeats the previous turn (prompt + answer) and adds to it finds the answer's blocks in
    the cache as well: finished sequences register the full blocks of everything whose keys / values they wrote."""
    from llmss_b200.utils.checkpoint import write_random_checkpoint

    path = str(write_random_checkpoint(tmp_path, "gptj-tiny", dtype=torch.float32, device="cpu", std=0.08))
    g = torch.Generator().manual_seed(21)
    first = torch.randint(1, 500, (40,), generator=g).tolist()
    kw = dict(device="cpu", max_batch=2, max_seq_len=160, block_size=16)
    plain, eng = Engine.from_pretrained(path, **kw), Engine.from_pretrained(path, prefix_cache=True, **kw)
    a1 = eng.generate([first], SamplingParams(max_new_tokens=30, is_greedy=True))[0]
    assert a1 == plain.generate([first], SamplingParams(max_new_tokens=30, is_greedy=True))[0]
    follow = first + a1 + torch.randint(1, 500, (5,), generator=g).tolist()          # 75 tokens: 69 of them were cached
    hits0 = eng.runner.stats.get("prefix_hit_tokens", 0)
    a2 = eng.generate([follow], SamplingParams(max_new_tokens=10, is_greedy=True))[0]
    assert eng.runner.stats["prefix_hit_tokens"] - hits0 == 64                      # 4 full blocks: prompt AND answer
    assert a2 == plain.generate([follow], SamplingParams(max_new_tokens=10, is_greedy=True))[0]
    a = eng.runner.alloc
    assert a.num_free == a.num_blocks - 1 and all(x == 0 for x in a.ref)
