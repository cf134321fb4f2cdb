"""Builds libbcg_b200.so (sm_100a only) in-tree with nvcc.

The shared library is the drop-in boundary (include/bcg_b200.h).  It is compiled into
byzantine_consensus_llm_agents_b200/_lib/ so that it travels with the source tree; there is
no JIT cache and no CPU fallback — if the library is missing the package refuses to run.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
OUT = PKG / "_lib"
LIB = OUT / "libbcg_b200.so"
LIB_TIMELINE = OUT / "libbcg_b200_tl.so"

SOURCES = [
    "capi.cu",
    "gemm_tcgen05.cu",
    "gemm2sm_tcgen05.cu",
    "norm_rope.cu",
    "attn_decode.cu",
    "attn_prefill.cu",
    "sampler.cu",
    "model_step.cu",
    "tp_comm.cu",
    "tp_reduce_norm.cu",
]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def _digest() -> str:
    h = hashlib.sha256()
    for p in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h"))
                    + [PKG.parent / "include" / "bcg_b200.h"]):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _build(lib: Path, defines, tag: str, force: bool, verbose: bool) -> Path:
    OUT.mkdir(exist_ok=True)
    stamp = OUT / f"build{tag}.stamp"
    digest = _digest() + " ".join(defines)
    if not force and lib.exists() and stamp.exists() and stamp.read_text().strip() == digest:
        return lib
    nvcc = _nvcc()
    extra = (["-Xptxas", "-v"] if verbose else []) + list(defines)

    def compile_one(src: str) -> Path:
        obj = OUT / (src + tag + ".o")
        cmd = [nvcc, *NVCC_FLAGS, *extra, "-c", str(CSRC / src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or r.returncode != 0:
            sys.stderr.write(f"$ {' '.join(cmd)}\n{r.stdout}{r.stderr}\n")
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [nvcc, "-shared", "-cudart", "static", "-o", str(lib), *map(str, objs), "-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    stamp.write_text(digest)
    return lib


def build(force: bool = False, verbose: bool = False) -> Path:
    return _build(LIB, [], "", force, verbose)


def build_timeline(force: bool = False, verbose: bool = False) -> Path:
    """Dev-only variant with %globaltimer stamps compiled in (tools/timeline.py); never loaded
    by the package itself."""
    return _build(LIB_TIMELINE, ["-DBCG_TIMELINE"], ".tl", force, verbose)


if __name__ == "__main__":
    fn = build_timeline if "--timeline" in sys.argv else build
    print(fn(force="--force" in sys.argv, verbose="-v" in sys.argv))
