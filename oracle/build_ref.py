"""Recipe that compiles the reference's CPU attention kernels, UNMODIFIED and where they
lie under /root/reference, into oracle/_ref/libvllm_cpu_attn_ref.so.

TEST INFRASTRUCTURE ONLY (checker + CPU baseline).  Inputs: csrc/cpu/cpu_attn.cpp,
csrc/cpu/utils.cpp and the headers they include (cpu_attn_impl.hpp, cpu_attn_vec*.hpp,
cpu_attn_amx.hpp, cpu_attn_fp8.hpp, cpu_types_x86.hpp ...), the dispatch header produced
by the reference's own generator csrc/cpu/generate_cpu_attn_dispatch.py (written to
oracle/_ref/, never into the read-only tree), and oracle/ref_cpu_attn_shim.cpp (ours:
only TORCH_LIBRARY registrations of the reference's entry points).  No reference source
is copied into this repository.  The full CPU build of the reference (oneDNN, numa) is
not attempted; see DESIGN.md.

On a machine without /root/reference (the GPU box) this is a no-op: the prebuilt .so
travels with the snapshot.
"""
from __future__ import annotations

import importlib.util
import os
import subprocess
import sys
import sysconfig
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF = Path(os.environ.get("VLLM_REFERENCE", "/root/reference"))
OUT_DIR = HERE / "_ref"
LIB = OUT_DIR / "libvllm_cpu_attn_ref.so"


def cpu_flags() -> set[str]:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def build(force: bool = False, verbose: bool = True) -> Path | None:
    if LIB.exists() and not force:
        if verbose:
            print(f"[oracle.build_ref] present: {LIB}")
        return LIB
    src_dir = REF / "csrc" / "cpu"
    if not (src_dir / "cpu_attn.cpp").exists():
        if verbose:
            print("[oracle.build_ref] /root/reference not present; nothing to build")
        return None
    flags = cpu_flags()
    if "avx512f" not in flags:
        if verbose:
            print("[oracle.build_ref] host CPU lacks AVX-512; reference CPU attention not built")
        return None
    OUT_DIR.mkdir(exist_ok=True)
    # 1) dispatch header, produced by the reference's own generator
    spec = importlib.util.spec_from_file_location("_ref_gen", src_dir / "generate_cpu_attn_dispatch.py")
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    (OUT_DIR / "cpu_attn_dispatch_generated.h").write_text(gen.generate_header_file())
    # 2) compile
    import torch

    tdir = Path(torch.__file__).resolve().parent
    inc, lib = tdir / "include", tdir / "lib"
    isa = ["-mavx512f", "-mavx512vl", "-mavx512bw", "-mavx512dq", "-mf16c", "-mfma"]
    defs = ["-DVLLM_NUMA_DISABLED", "-DVLLM_CPU_EXTENSION", "-D_GLIBCXX_USE_CXX11_ABI=1"]
    if {"amx_bf16", "amx_tile", "avx512_bf16"} <= flags:
        isa += ["-mamx-bf16", "-mamx-tile", "-mavx512bf16", "-mavx512vnni"]
        defs += ["-DCPU_CAPABILITY_AMXBF16"]
    elif "avx512_bf16" in flags:
        isa += ["-mavx512bf16", "-mavx512vnni"]
    cmd = [
        "g++", "-O3", "-std=c++17", "-fPIC", "-shared", "-fopenmp", *defs, *isa,
        f"-I{OUT_DIR}", f"-I{REF / 'csrc'}", f"-I{src_dir}", f"-I{inc}",
        f"-I{inc / 'torch' / 'csrc' / 'api' / 'include'}", f"-I{sysconfig.get_paths()['include']}",
        str(src_dir / "cpu_attn.cpp"), str(src_dir / "utils.cpp"), str(HERE / "ref_cpu_attn_shim.cpp"),
        f"-L{lib}", "-ltorch", "-ltorch_cpu", "-lc10", f"-Wl,-rpath,{lib}", "-o", str(LIB),
    ]
    if verbose:
        print("[oracle.build_ref] compiling reference CPU attention (several minutes) ...")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"reference CPU attention build failed:\n{r.stderr[-4000:]}")
    (OUT_DIR / "BUILD_INFO.txt").write_text("flags: " + " ".join(isa + defs) + "\n")
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv)
