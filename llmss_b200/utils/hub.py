"""Weight-file discovery.

Parity: reference `src/llmss/server/models/utils/hub.py:16-163` -- `weight_files`,
`weight_hub_files`, `try_to_load_from_cache`, `download_weights`, and the
`WEIGHTS_CACHE_OVERRIDE` / `HUGGINGFACE_HUB_CACHE` environment variables.

A local directory is resolved by globbing; a hub id is resolved through the local HF
cache layout (``models--org--name/{refs,snapshots}``).  Network calls are only made
by `weight_hub_files` / `download_weights` and are imported lazily so the serving
path has no hard dependency on `huggingface_hub`.
"""
from __future__ import annotations

import os
import time
from datetime import timedelta
from pathlib import Path
from typing import List, Optional

from .logging import logger

WEIGHTS_CACHE_OVERRIDE = os.getenv("WEIGHTS_CACHE_OVERRIDE", None)


def _hub_cache_dir() -> Path:
    env = os.getenv("HUGGINGFACE_HUB_CACHE")
    if env:
        return Path(env)
    try:
        from huggingface_hub.constants import HUGGINGFACE_HUB_CACHE

        return Path(HUGGINGFACE_HUB_CACHE)
    except Exception:
        return Path.home() / ".cache" / "huggingface" / "hub"


class EntryNotFound(FileNotFoundError):
    """Raised when a model/revision has no file with the wanted extension."""


_SKIP_WORDS = ("arguments", "args", "training")


def weight_hub_files(model_id: str, revision: Optional[str] = None, extension: str = ".safetensors") -> List[str]:
    """Top-level weight filenames of a hub repo (ref hub.py:19-39). Needs network."""
    from huggingface_hub import HfApi

    info = HfApi().model_info(model_id, revision=revision)
    names = [
        s.rfilename
        for s in info.siblings
        if s.rfilename.endswith(extension)
        and "/" not in s.rfilename
        and not any(w in s.rfilename for w in _SKIP_WORDS)
    ]
    if not names:
        raise EntryNotFound(f"No {extension} weights found for model {model_id} and revision {revision}.")
    return names


def try_to_load_from_cache(model_id: str, revision: Optional[str], filename: str) -> Optional[Path]:
    """Look a file up in the local HF cache without touching the network (ref hub.py:42-74)."""
    revision = revision or "main"
    repo_cache = _hub_cache_dir() / f"models--{model_id.replace('/', '--')}"
    if not repo_cache.is_dir():
        return None
    ref_file = repo_cache / "refs" / revision
    if ref_file.is_file():
        revision = ref_file.read_text().strip()
    snapshots = repo_cache / "snapshots"
    if not snapshots.is_dir() or revision not in os.listdir(snapshots):
        return None
    cached = snapshots / revision / filename
    return cached if cached.is_file() else None


def _cached_snapshot_files(model_id: str, revision: Optional[str], extension: str) -> List[Path]:
    """All cached files with `extension` for a hub id (offline replacement for the hub listing)."""
    revision = revision or "main"
    repo_cache = _hub_cache_dir() / f"models--{model_id.replace('/', '--')}"
    ref_file = repo_cache / "refs" / revision
    if ref_file.is_file():
        revision = ref_file.read_text().strip()
    snap = repo_cache / "snapshots" / revision
    if not snap.is_dir():
        return []
    return sorted(p for p in snap.glob(f"*{extension}") if not any(w in p.name for w in _SKIP_WORDS))


def weight_files(model_id: str, revision: Optional[str] = None, extension: str = ".safetensors") -> List[Path]:
    """Local weight files for a directory or (cached) hub id (ref hub.py:77-118)."""
    p = Path(model_id)
    if p.exists() and p.is_dir():
        local = sorted(p.glob(f"*{extension}"))
        if not local:
            raise FileNotFoundError(f"No local weights found in {model_id} with extension {extension}")
        return local

    try:
        filenames = weight_hub_files(model_id, revision, extension)
    except EntryNotFound:
        if extension != ".safetensors":
            raise
        # .bin on the hub may have been converted locally: map names (ref hub.py:89-96)
        pt = weight_hub_files(model_id, revision, extension=".bin")
        filenames = [f"{Path(f).stem.removeprefix('pytorch_')}.safetensors" for f in pt]
    except Exception:
        # No network: fall back to whatever the local cache holds for this id.
        cached = _cached_snapshot_files(model_id, revision, extension)
        if cached and WEIGHTS_CACHE_OVERRIDE is None:
            return cached
        filenames = [c.name for c in cached]
        if not filenames:
            raise FileNotFoundError(
                f"Model {model_id} is neither a local directory nor present in the local HF cache "
                f"({_hub_cache_dir()}), and the hub is unreachable."
            )

    override = os.getenv("WEIGHTS_CACHE_OVERRIDE", WEIGHTS_CACHE_OVERRIDE)
    if override is not None:
        files = []
        for name in filenames:
            fp = Path(override) / name
            if not fp.exists():
                raise FileNotFoundError(f"File {fp} not found in {override}.")
            files.append(fp)
        return files

    files = []
    for name in filenames:
        cached = try_to_load_from_cache(model_id, revision=revision, filename=name)
        if cached is None:
            raise FileNotFoundError(
                f"File {name} of model {model_id} not found in {_hub_cache_dir()}. "
                f"Download the weights first (llmss_b200.utils.hub.download_weights)."
            )
        files.append(cached)
    return files


def download_weights(filenames: List[str], model_id: str, revision: Optional[str] = None,
                     tries: int = 5, backoff: float = 5.0) -> List[Path]:
    """Fetch files into the HF cache with retry/back-off and ETA logging (ref hub.py:121-163)."""
    from huggingface_hub import hf_hub_download

    def fetch(name: str) -> Path:
        hit = try_to_load_from_cache(model_id, revision, name)
        if hit is not None:
            logger.info(f"File {name} already present in cache.")
            return hit
        for attempt in range(tries):
            try:
                t0 = time.time()
                local = hf_hub_download(filename=name, repo_id=model_id, revision=revision, local_files_only=False)
                logger.info(f"Downloaded {local} in {timedelta(seconds=int(time.time() - t0))}.")
                return Path(local)
            except Exception as e:  # noqa: BLE001
                if attempt + 1 == tries:
                    raise
                logger.error(str(e))
                logger.info(f"Retrying in {backoff} seconds ({attempt + 1}/{tries - 1})")
                time.sleep(backoff)
        raise RuntimeError("unreachable")

    t_start = time.time()
    out = []
    for i, name in enumerate(filenames):
        out.append(fetch(name))
        elapsed = timedelta(seconds=int(time.time() - t_start))
        left = len(filenames) - (i + 1)
        eta = (elapsed / (i + 1)) * left if left else 0
        logger.info(f"Download: [{i + 1}/{len(filenames)}] -- ETA: {eta}")
    return out
