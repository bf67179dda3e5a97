"""Build the sm_100a kernel library in-tree with nvcc (no GPU needed: cross-compiles).

``python -m llmss_b200.ops.build`` or ``build_library()``.  Objects are cached under
``build/obj`` keyed by a hash of the source, its local includes and the flags, and the final
``llmss_b200/_C/libllmss_b200.so`` is what ``llmss_b200.ops._lib`` dlopens (ctypes).  The
.so is git-ignored but travels with the working tree to the GPU box.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
OBJ_DIR = ROOT / "build" / "obj"
LIB_DIR = ROOT / "llmss_b200" / "_C"
LIB_PATH = LIB_DIR / "libllmss_b200.so"

BASE_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    "--expt-relaxed-constexpr",
]
# include paths are location dependent and must not enter the staleness digest: the tree (with its
# prebuilt .so) is copied to other machines under a different root.
NVCC_FLAGS = BASE_FLAGS + ["-I", str(ROOT / "include"), "-I", str(CSRC)]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def sources():
    return sorted(CSRC.glob("*.cu"))


def _digest(src: Path) -> str:
    h = hashlib.sha256()
    h.update(" ".join(BASE_FLAGS).encode())
    h.update(src.read_bytes())
    for inc in sorted(list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))):
        h.update(inc.read_bytes())
    return h.hexdigest()[:16]


def _compile(src: Path, verbose: bool) -> Path:
    obj = OBJ_DIR / f"{src.stem}.{_digest(src)}.o"
    if obj.exists():
        return obj
    for old in OBJ_DIR.glob(f"{src.stem}.*.o"):
        old.unlink()
    cmd = [_nvcc(), *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
    if verbose:
        print("[build]", " ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    return obj


def tree_digest() -> str:
    """Digest of everything that determines the library contents."""
    return hashlib.sha256("".join(_digest(s) for s in sources()).encode()).hexdigest()[:24]


def is_current() -> bool:
    stamp = LIB_DIR / "build.stamp"
    return LIB_PATH.exists() and stamp.exists() and stamp.read_text().strip() == tree_digest()


def build_library(verbose: bool = False, force: bool = False) -> Path:
    if not force and is_current():  # e.g. on the GPU box: the .so travelled with the tree
        return LIB_PATH
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    # one builder at a time (torchrun starts N ranks at once); the others wait, then see a current library
    import fcntl

    with open(LIB_DIR / ".build.lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and is_current():
                return LIB_PATH
            return _build_locked(verbose, force)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(verbose: bool, force: bool) -> Path:
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    if force:
        for o in OBJ_DIR.glob("*.o"):
            o.unlink()
    srcs = sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), srcs))
    stamp = LIB_DIR / "build.stamp"
    want = tree_digest()
    tmp = LIB_PATH.with_suffix(f".so.tmp{os.getpid()}")
    cmd = [_nvcc(), "-shared", "-o", str(tmp), *map(str, objs)]
    if verbose:
        print("[build]", " ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, LIB_PATH)  # atomic: a concurrent dlopen never sees a half-written file
    stamp.write_text(want)
    return LIB_PATH


if __name__ == "__main__":
    p = build_library(verbose=True, force="--force" in sys.argv)
    print(f"built {p} ({p.stat().st_size / 1e6:.1f} MB)")
