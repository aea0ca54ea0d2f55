"""Loader + thin callers for oracle/_ref/libvllm_cpu_attn_ref.so — the reference's own
CPU attention kernels (csrc/cpu/cpu_attn.cpp) compiled unmodified by oracle/build_ref.py.

TEST INFRASTRUCTURE ONLY: used to pin oracle.py and as bench.py's "reference" CPU
baseline.  Call pattern follows vllm/v1/attention/backends/cpu_attn.py:225-275,336-437
(get_scheduler_metadata -> cpu_attn_reshape_and_cache -> cpu_attention_with_kv_cache;
cache layout [num_blocks, num_kv_heads, block_size, head_size]).
"""
from __future__ import annotations

from pathlib import Path

import torch

from .build_ref import LIB, cpu_flags

_loaded = False


def available() -> bool:
    """The .so exists AND this host can execute it (it is built with AVX-512 / AMX)."""
    if not LIB.exists():
        return False
    flags = cpu_flags()
    info = LIB.parent / "BUILD_INFO.txt"
    need = {"avx512f", "avx512bw", "avx512vl", "avx512dq"}
    if info.exists() and "amx" in info.read_text():
        need |= {"amx_bf16", "amx_tile", "avx512_bf16"}
    elif not info.exists():
        need |= {"amx_bf16", "amx_tile", "avx512_bf16"}  # first build in this repo used AMX
    return need <= flags


def load() -> None:
    global _loaded
    if not _loaded:
        if not available():
            raise RuntimeError("reference CPU attention library is not available on this host")
        torch.ops.load_library(str(LIB))
        _loaded = True


def pick_isa(dtype: torch.dtype, block_size: int, head_size: int) -> str:
    """vllm/v1/attention/backends/cpu_attn.py:480-519 (_get_attn_isa), x86 branch."""
    if head_size % 32 != 0 and head_size % 16 == 0:
        return "vec16"
    flags = cpu_flags()
    if "amx_tile" in flags and dtype == torch.bfloat16 and block_size % 32 == 0:
        return "amx"
    return "vec" if block_size % 32 == 0 else "vec16"


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, isa, k_scale=1.0, v_scale=1.0,
                      kv_cache_dtype="auto") -> None:
    load()
    torch.ops.vllm_cpu_ref.cpu_attn_reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, isa,
                                                      k_scale, v_scale, kv_cache_dtype)


def decode_attention(query, key_cache, value_cache, block_table, seq_lens, scale, isa, k_scale=1.0, v_scale=1.0,
                     kv_cache_dtype="auto", sliding_window=-1, softcap=0.0):
    """query [S, Hq, D] (one token per sequence); caches [num_blocks, Hkv, block_size, D]."""
    load()
    S, Hq, D = query.shape
    Hkv = key_cache.size(1)
    qsl = torch.arange(S + 1, dtype=torch.int32)
    md = torch.ops.vllm_cpu_ref.get_scheduler_metadata(S, Hq, Hkv, D, seq_lens, query.dtype, qsl, True,
                                                       sliding_window, isa, False, None, kv_cache_dtype)
    out = torch.empty_like(query)
    torch.ops.vllm_cpu_ref.cpu_attention_with_kv_cache(query, key_cache, value_cache, out, qsl, seq_lens, scale,
                                                       True, None, sliding_window, block_table, softcap, md, None,
                                                       None, k_scale, v_scale, kv_cache_dtype)
    return out
