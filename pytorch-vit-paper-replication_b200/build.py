"""Build libb200vit.so (hand-written sm_100a kernels + C ABI) in-tree with nvcc.

    python pytorch-vit-paper-replication_b200/build.py [--force] [--verbose]

nvcc cross-compiles for sm_100a without a GPU. The library is written to
pytorch-vit-paper-replication_b200/b200vit/lib/libb200vit.so (git-ignored, travels with gpurun).
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
OUT_DIR = HERE / "b200vit" / "lib"
OBJ_DIR = HERE / "build"
LIB = OUT_DIR / "libb200vit.so"

SOURCES = [
    "runtime.cu",
    "api.cu",
    "gemm_sm100.cu",
    "attention.cu",
    "attention_tc.cu",
    "layernorm.cu",
    "elementwise.cu",
    "head_loss.cu",
    "optim.cu",
    "peer_reduce.cu",
    "fp32_mode.cu",
    "block.cu",
]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _digest() -> str:
    h = hashlib.sha256()
    for p in sorted(CSRC.glob("*")) + [HERE.parent / "include" / "b200vit.h", Path(__file__)]:
        if p.is_file():
            h.update(p.name.encode())
            h.update(p.read_bytes())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> Path:
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    stamp = OBJ_DIR / "digest.txt"
    digest = _digest()
    if not force and LIB.exists() and stamp.exists() and stamp.read_text() == digest:
        return LIB
    nvcc = _nvcc()
    sources = [s for s in SOURCES if (CSRC / s).exists()]

    def compile_one(src: str):
        obj = OBJ_DIR / (src.replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, "-c", str(CSRC / src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        return src, obj, r

    objs = []
    with cf.ThreadPoolExecutor(max_workers=min(8, len(sources))) as ex:
        for src, obj, r in ex.map(compile_one, sources):
            log = (r.stdout or "") + (r.stderr or "")
            (OBJ_DIR / (src + ".log")).write_text(log)
            if r.returncode != 0:
                sys.stderr.write(log)
                raise RuntimeError(f"nvcc failed on {src}")
            if verbose:
                sys.stderr.write(log)
            objs.append(str(obj))
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB), *objs, "-lcudart_static",
           "-ldl", "-lpthread", "-lrt"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    stamp.write_text(digest)
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(force=a.force, verbose=a.verbose))
