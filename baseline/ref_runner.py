"""Runs the UNMODIFIED reference package (installed with pip --target baseline/_ref) so that
bench.py --impl reference measures the reference's own code path on this box.

Install (recorded in DESIGN.md):
    python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
        --target baseline/_ref /root/reference

Three in-process shims are needed because the reference pins torch 2.0.1 / transformers 4.30.2 /
accelerate 0.21.0 and this image has torch 2.11 / transformers 5.5 / no accelerate (SURVEY.md
section 6.3).  None of them touches the reference's files or its compute path:
  1. a stub `accelerate` module whose `init_empty_weights` is `torch.device("meta")`
     (used only while constructing LayerNorm shells, reference layers.py:6,16,27);
  2. `PreTrainedModel.get_head_mask` (removed in transformers 5) returning `[None] * n`;
  3. `config._attn_implementation = "eager"` so transformers 5 accepts GPTBigCode construction.
"""
from __future__ import annotations

import contextlib
import os
import sys
import types
from pathlib import Path

import torch

REF_ROOT = Path(__file__).resolve().parent / "_ref"


def _install_shims():
    import transformers  # must be imported before the accelerate stub (it probes accelerate.__spec__)
    from transformers import PreTrainedModel

    if "accelerate" not in sys.modules:
        try:
            import accelerate  # noqa: F401
        except Exception:
            acc = types.ModuleType("accelerate")

            @contextlib.contextmanager
            def init_empty_weights(include_buffers: bool = False):
                with torch.device("meta"):
                    yield

            acc.init_empty_weights = init_empty_weights
            acc.__spec__ = None
            sys.modules["accelerate"] = acc
    if not hasattr(PreTrainedModel, "get_head_mask"):
        PreTrainedModel.get_head_mask = lambda self, head_mask, n, *a, **k: [None] * n
    try:
        import loguru  # noqa: F401
    except Exception:
        lg = types.ModuleType("loguru")
        import logging

        lg.logger = logging.getLogger("llmss-ref")
        sys.modules["loguru"] = lg
    return transformers


def import_reference():
    if not (REF_ROOT / "llmss").is_dir():
        raise RuntimeError(f"{REF_ROOT}/llmss not found: run the pip --target install described in DESIGN.md")
    _install_shims()
    if str(REF_ROOT) not in sys.path:
        sys.path.insert(0, str(REF_ROOT))
    from llmss.server.models.custom_modeling import MODEL_REGISTRY
    from llmss.server.models.utils.dist import initialize_torch_distributed
    from llmss.server.models.utils.hub import weight_files
    from llmss.server.models.utils.weights import Weights

    return MODEL_REGISTRY, initialize_torch_distributed, weight_files, Weights


_PG = None


def init_reference_dist():
    """The reference's own process-group bring-up (dist.py:40-77); idempotent."""
    global _PG
    if _PG is None:
        _, initialize_torch_distributed, _, _ = import_reference()
        _PG = initialize_torch_distributed()
    return _PG


def build_reference_model(path: str, local_rank: int, dtype=torch.float16, device=None):
    """Same bring-up sequence as the reference's generate.py:42-68."""
    from transformers import AutoConfig

    MODEL_REGISTRY, _, weight_files, Weights = import_reference()
    process_group, rank, world_size = init_reference_dist()
    if device is None:
        device = torch.device(f"cuda:{local_rank}") if torch.cuda.is_available() else torch.device("cpu")
    config = AutoConfig.from_pretrained(path)
    config._attn_implementation = "eager"
    filenames = weight_files(path)
    weights = Weights(filenames, device=device, dtype=dtype, process_group=process_group)
    model = MODEL_REGISTRY[config.model_type](config, weights)
    model.eval()
    return model, process_group, config
