"""ctypes binding of libb200vllm.so (the C ABI declared in include/b200_vllm.h).

There is NO fallback: if the library is missing or an entry point fails, a
RuntimeError is raised (the reference raises from STD_TORCH_CHECK the same way).
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("B200_VLLM_LIB", _PKG / "libb200vllm.so"))

# error codes (include/b200_vllm.h)
OK, ERR_INVALID_ARG, ERR_ALIGNMENT, ERR_UNSUPPORTED, ERR_CUDA, ERR_WORKSPACE = 0, -1, -2, -3, -4, -5
DT_F32, DT_F16, DT_BF16 = 0, 1, 2
KV_AUTO, KV_FP8_E4M3, KV_FP8_E5M2 = 0, 1, 2

i32, i64, f32, vp, sz = C.c_int, C.c_int64, C.c_float, C.c_void_p, C.c_size_t


class PagedDecodeParams(C.Structure):
    """struct b200_paged_decode_params (include/b200_vllm.h)."""

    _fields_ = [
        ("out", vp), ("query", vp), ("key_cache", vp), ("value_cache", vp),
        ("block_table", vp), ("seq_lens", vp), ("k_scale", vp), ("v_scale", vp),
        ("out_lse", vp), ("token_to_seq", vp), ("workspace", vp), ("workspace_bytes", sz),
        ("q_stride_seq", i64), ("q_stride_head", i64),
        ("out_stride_seq", i64), ("out_stride_head", i64),
        ("kv_block_stride", i64), ("kv_page_stride", i64), ("kv_head_stride", i64),
        ("v_block_stride", i64), ("v_page_stride", i64), ("v_head_stride", i64),
        ("block_table_stride", i64), ("num_blocks", i64),
        ("num_seqs", i32), ("num_q_heads", i32), ("num_kv_heads", i32),
        ("head_size", i32), ("block_size", i32), ("max_seq_len", i32),
        ("q_dtype", i32), ("kv_dtype", i32), ("num_splits", i32),
        ("scale", f32), ("logits_soft_cap", f32), ("sliding_window", i32),
        ("output_scale", vp),
    ]


# name -> (restype, argtypes); must list every symbol of include/b200_vllm.h
SIGNATURES: dict[str, tuple] = {
    "b200_version": (C.c_char_p, []),
    "b200_last_error": (C.c_char_p, []),
    "b200_num_kernels_launched": (i32, []),
    "b200_reshape_and_cache_flash": (i32, [vp, vp, vp, vp, vp, i64, i32, i32, i32, i64, i64, i64, i64, i64,
                                           i32, i32, vp, vp, i32, vp]),
    "b200_compute_slot_mapping": (i32, [vp, i64, i64, vp, i32, vp, vp, i64, i32, vp]),
    "b200_paged_decode_workspace_bytes": (sz, [i32, i32, i32, i32]),
    "b200_paged_decode_pick_splits": (i32, [i32, i32, i32, i32]),
    "b200_paged_decode_pick_splits_ex": (i32, [i32, i32, i32, i32, i32, i32]),
    "b200_paged_attention_decode": (i32, [C.POINTER(PagedDecodeParams), vp]),
    "b200_paged_decode_static_scratch_bytes": (sz, []),
    "b200_paged_attention_decode_static_scratch": (i32, [C.POINTER(PagedDecodeParams), vp]),
    "b200_merge_attn_states": (i32, [vp, vp, vp, vp, vp, vp, i64, i32, i32, i64, i64, i64, i64, i64, i64, i64,
                                     i32, vp]),
    "b200_static_scaled_fp8_quant": (i32, [vp, vp, vp, i64, i32, i64, i64, i32, i32, i64, i64, i32, vp]),
    "b200_dynamic_scaled_fp8_quant": (i32, [vp, vp, vp, i64, i32, i64, i64, i32, vp]),
    "b200_dynamic_per_token_scaled_fp8_quant": (i32, [vp, vp, vp, vp, i64, i32, i64, i64, i32, vp]),
    "b200_scaled_fp4_quant": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]),
    "b200_scaled_mm_fp8": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, i64, i64, i64, i32, i32, i32, vp, sz, vp]),
    "b200_scaled_mm_workspace_bytes": (sz, [i32, i32]),
    "b200_scaled_mm_nvfp4": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, i64, i32, vp]),
    "b200_rms_norm": (i32, [vp, vp, vp, vp, i64, i32, i64, f32, i32, vp]),
    "b200_silu_and_mul": (i32, [vp, vp, i64, i32, i32, vp]),
    "b200_rotary_embedding_neox": (i32, [vp, vp, vp, vp, i64, i32, i32, i32, i32, i64, i64, i32, vp]),
    "b200_rms_norm_dynamic_per_token_quant": (i32, [vp, vp, vp, vp, f32, vp, vp, i64, i32, i64, i32, vp]),
    "b200_silu_and_mul_dynamic_per_token_quant": (i32, [vp, vp, vp, i64, i32, i32, vp]),
    "b200_rotary_embedding_and_cache_write": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, i64, i64, i32, i32, i32, i32,
                                                    i64, i64, i64, i32, i64, i64, i64, i32, i32, vp, vp, i32, vp]),
    "b200_peer_all_reduce": (i32, [vp, vp, i32, i32, i64, i32, i64, i64, vp]),
    "b200_peer_all_reduce_out": (i32, [vp, vp, vp, i32, i32, i64, i32, i64, i64, vp]),
    "b200_peer_all_reduce_rms_norm_quant": (i32, [vp, vp, vp, vp, i32, i32, i64, i64, vp, vp, f32, i64, i32, i32, vp]),
    "b200_nvls_all_reduce": (i32, [C.c_uint64, vp, vp, i32, i32, i64, i32, i64, i64, vp]),
    "b200_silu_and_mul_nvfp4_quant": (i32, [vp, vp, vp, vp, i32, i32, i32, vp]),
    "b200_rms_norm_nvfp4_quant": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i64, f32, i32, vp]),
    "b200_lamport_workspace_bytes": (i64, [i32, i64]),
    "b200_lamport_workspace_init": (i32, [vp, i64, vp]),
    "b200_lamport_all_reduce": (i32, [vp, vp, vp, i64, i32, i32, i64, i32, i32, vp]),
    "b200_lamport_all_reduce_rms_norm_quant": (i32, [vp, vp, vp, vp, i64, i32, i32, vp, vp, f32, i64, i32, i32, vp]),
    "b200_lamport_all_reduce_rms_norm_nvfp4_quant": (i32, [vp, vp, vp, vp, i64, i32, i32, vp, vp, vp, f32, i64, i32, i32, vp]),
    "b200_nvls_all_reduce_out": (i32, [vp, C.c_uint64, vp, vp, i32, i32, i64, i32, i64, i64, vp]),
    "b200_nvls_all_reduce_rms_norm_quant": (i32, [vp, vp, C.c_uint64, vp, vp, i32, i32, i64, i64, vp, vp, f32, i64, i32,
                                                  i32, vp]),
    "b200_peer_set_barrier_timeout": (None, [C.c_ulonglong]),
    "b200_set_pdl": (None, [i32]),
    "b200_get_pdl": (i32, []),
    "b200_debug_gemm_trace": (None, [vp]),
    "b200_debug_attn_trace": (None, [vp]),
    "b200_debug_attn_fp8_tcgen05": (None, [i32]),
    "b200_debug_gemm_mode": (None, [i32]),
    "b200_debug_gemm_dump": (None, [vp]),
}

_lib: C.CDLL | None = None


def load() -> C.CDLL:
    """Load the shared library (once). Raises RuntimeError when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RuntimeError(
            f"{LIB_PATH} not found: the CUDA extension is not built. Run `python -m vllm_b200.build` "
            "(or __graft_entry__.build()). vllm_b200 has no CPU or library fallback."
        )
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError -> missing export, fail loudly
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def is_loaded() -> bool:
    return _lib is not None


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().b200_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def num_kernels_launched() -> int:
    return int(load().b200_num_kernels_launched())
