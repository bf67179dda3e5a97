"""Reference import paths (compat) and the generate.py CLI contract on CPU, TP=1 (BASELINE config 1)."""
import json
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parents[1]


def test_compat_import_paths():
    code = ("import llmss_b200.compat\n"
            "from llmss.server.models.custom_modeling import MODEL_REGISTRY\n"
            "from llmss.server.models.utils.dist import initialize_torch_distributed, FakeGroup\n"
            "from llmss.server.models.utils.hub import weight_files\n"
            "from llmss.server.models.utils.weights import Weights\n"
            "from llmss.server.models.utils.layers import TensorParallelRowLinear, TensorParallelColumnLinear, "
            "TensorParallelEmbedding, TensorParallelHead, FastLinear, SuperLayer, get_linear\n"
            "print(sorted(MODEL_REGISTRY))\n")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=str(ROOT), timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "gptj" in r.stdout and "gpt_bigcode" in r.stdout


def _make_gpt2_dir(tmp_path):
    """GPT-2-small-shaped tiny model + a byte-level tokenizer built offline."""
    from tokenizers import Tokenizer, models, pre_tokenizers, decoders
    from transformers import PreTrainedTokenizerFast

    from llmss_b200.utils.checkpoint import write_random_checkpoint

    cfg = dict(model_type="gpt2", architectures=["GPT2LMHeadModel"], vocab_size=260, n_positions=128, n_embd=64,
               n_layer=2, n_head=4, activation_function="gelu_new", layer_norm_epsilon=1e-5, tie_word_embeddings=True,
               bos_token_id=256, eos_token_id=256)
    path = write_random_checkpoint(tmp_path, cfg, dtype=torch.float32, device="cpu", std=0.1)
    vocab = {chr(0x100 + i) if i < 33 or i > 126 else chr(i): i for i in range(256)}
    vocab.update({"<|endoftext|>": 256, "<pad>": 257})
    # a character-level "BPE" with no merges over a printable alphabet
    vocab = {c: i for i, c in enumerate([chr(i) for i in range(32, 127)])}
    vocab.update({"<|endoftext|>": len(vocab), "<pad>": len(vocab) + 1})
    tok = Tokenizer(models.BPE(vocab=vocab, merges=[], unk_token="<|endoftext|>"))
    tok.pre_tokenizer = pre_tokenizers.Split("", "isolated")
    tok.decoder = decoders.Fuse()
    fast = PreTrainedTokenizerFast(tokenizer_object=tok, eos_token="<|endoftext|>", pad_token="<pad>")
    fast.save_pretrained(str(path))
    return path


def test_generate_cli_cpu_tp1(tmp_path):
    path = _make_gpt2_dir(tmp_path)
    cmd = [sys.executable, str(ROOT / "generate.py"), "--pretrained_model_path", str(path), "--prompts", "hello world",
           "the quick brown fox", "--max_new_tokens", "6", "--is_greedy", "--use_cache", "--dtype", "float32"]
    env = {"PATH": "/usr/bin:/bin", "WORLD_SIZE": "1", "RANK": "0", "CUDA_VISIBLE_DEVICES": ""}
    import os

    r = subprocess.run(cmd, capture_output=True, text=True, cwd=str(ROOT), timeout=600, env={**os.environ, **env})
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert lines[-3].startswith("elapsed time: ")
    assert lines[-2] == "prompts: ['hello world', 'the quick brown fox']"
    assert lines[-1].startswith("continuations: [") and lines[-1].count("'") >= 4
    # determinism + flag validation
    r2 = subprocess.run(cmd, capture_output=True, text=True, cwd=str(ROOT), timeout=600, env={**os.environ, **env})
    assert r2.stdout.splitlines()[-1] == lines[-1]
    bad = subprocess.run(cmd[:-2] + ["--temperature", "2.0"], capture_output=True, text=True, cwd=str(ROOT), timeout=600,
                         env={**os.environ, **env})
    assert bad.returncode != 0 and "temperature" in bad.stderr
