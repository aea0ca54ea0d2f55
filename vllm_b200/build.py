"""Build libb200vllm.so (sm_100a only) in-tree with plain nvcc.

`python -m vllm_b200.build` or `build()` from `__graft_entry__`.  nvcc
cross-compiles without a GPU; the .so is git-ignored but travels with the
repo snapshot to the GPU box.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shlex
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
OBJ_DIR = PKG_DIR / "_build"
LIB_PATH = PKG_DIR / "libb200vllm.so"

SOURCES = [
    "runtime.cu",
    "cache_write.cu",
    "paged_decode.cu",
    "paged_decode_fp8.cu",
    "fp8_quant.cu",
    "nvfp4_quant.cu",
    "merge_elementwise.cu",
    "fused_quant.cu",
    "fp8_gemm.cu",
    "fp8_gemm_decode.cu",
    "gemm_persist.cu",
    "nvfp4_gemm.cu",
    "peer_allreduce.cu",
    "lamport_allreduce.cu",
]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-std=c++17", "-O3", "-lineinfo",
    "--expt-relaxed-constexpr", "--extended-lambda",
    "-Xcompiler", "-fPIC",
    *shlex.split(os.environ.get("B200_NVCC_EXTRA", "")),  # experiments: -DB200_EXP_... switches
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; libb200vllm.so cannot be built")


def _digest(paths: list[Path]) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _compile_one(nvcc: str, src: Path, obj: Path) -> str:
    cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    return src.name


def build(force: bool = False, verbose: bool = True) -> Path:
    """Compile every CUDA source for sm_100a and link libb200vllm.so."""
    nvcc = _nvcc()
    OBJ_DIR.mkdir(exist_ok=True)
    headers = list(CSRC.glob("*.cuh")) + [PKG_DIR.parent / "include" / "b200_vllm.h"]
    srcs = [CSRC / s for s in SOURCES]
    stamp = OBJ_DIR / "stamp.txt"
    digest = _digest(srcs + headers)
    if not force and LIB_PATH.exists() and stamp.exists() and stamp.read_text() == digest:
        if verbose:
            print(f"[vllm_b200.build] up to date: {LIB_PATH}")
        return LIB_PATH
    objs = [OBJ_DIR / (s.stem + ".o") for s in srcs]
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        futs = [ex.submit(_compile_one, nvcc, s, o) for s, o in zip(srcs, objs)]
        for f in cf.as_completed(futs):
            name = f.result()
            if verbose:
                print(f"[vllm_b200.build] compiled {name}")
    link = [nvcc, "-shared", "-o", str(LIB_PATH), *map(str, objs)]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(digest)
    if verbose:
        print(f"[vllm_b200.build] linked {LIB_PATH}")
    return LIB_PATH


if __name__ == "__main__":
    build(force="--force" in sys.argv)
