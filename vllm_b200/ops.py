"""Host-side operator layer: same names, argument meaning and error behaviour as
`vllm/_custom_ops.py` for the decode hot path, bottoming out in the C ABI.

Reference wrappers mirrored here (path:line under /root/reference):
  reshape_and_cache_flash      vllm/_custom_ops.py:2631-2650
  scaled_fp8_quant             vllm/_custom_ops.py:1832-1899
  cutlass_scaled_mm            vllm/_custom_ops.py:756-804
  scaled_fp4_quant             vllm/_custom_ops.py:1523-1598 (+ create_fp4_scale_tensor :36-70)
  cutlass_scaled_fp4_mm        vllm/_custom_ops.py:733-745
  merge_attn_states            vllm/_custom_ops.py:178-198
  paged attention decode       vllm/_custom_ops.py:120-163 (legacy v1/v2) /
                               vllm/v1/attention/ops/triton_unified_attention.py (current)

Everything runs on the CURRENT torch CUDA stream of the tensors' device, never
synchronises and never allocates behind the caller's back except through
torch's caching allocator (CUDA-graph capturable, like the reference).
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _lib
from ._lib import DT_BF16, DT_F16, DT_F32, KV_AUTO, KV_FP8_E4M3, KV_FP8_E5M2

FP8_DTYPE = torch.float8_e4m3fn
PAD_SLOT_ID = -1

_DT = {torch.float32: DT_F32, torch.float16: DT_F16, torch.bfloat16: DT_BF16}


def _dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise RuntimeError(f"vllm_b200: unsupported dtype {t.dtype}") from None


def kv_dtype_code(kv_cache_dtype: str) -> int:
    """'auto'/'bfloat16'/'float16' -> AUTO; 'fp8'/'fp8_e4m3' -> e4m3; 'fp8_e5m2' -> e5m2
    (csrc/attention/dtype_fp8.cuh:21; quant_utils.cuh:575-613)."""
    if kv_cache_dtype in ("auto", "bfloat16", "float16", "half"):
        return KV_AUTO
    if kv_cache_dtype in ("fp8", "fp8_e4m3"):
        return KV_FP8_E4M3
    if kv_cache_dtype == "fp8_e5m2":
        return KV_FP8_E5M2
    raise RuntimeError(f"Unsupported data type of kv cache: {kv_cache_dtype}")


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _require_cuda(*ts: torch.Tensor | None) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("vllm_b200 ops need CUDA tensors (sm_100a); there is no CPU path")


# --------------------------------------------------------------------------- KV write
def reshape_and_cache_flash(
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    """key/value [T, H, D]; caches are 4-D views (num_blocks, block_size, H, D) with any
    strides (NHD, HND, interleaved); num_tokens = slot_mapping.size(0) (may be < T)."""
    _require_cuda(key, value, key_cache, value_cache, slot_mapping)
    if key.dim() != 3 or value.dim() != 3 or key_cache.dim() != 4 or value_cache.dim() != 4:
        raise RuntimeError("reshape_and_cache_flash: key/value must be 3-D and caches 4-D")
    if key.stride(2) != 1 or value.stride(2) != 1 or key.stride(1) != key.size(2) or value.stride(1) != value.size(2):
        raise RuntimeError("reshape_and_cache_flash: key/value rows must be [H, D] contiguous")
    if key_cache.stride(3) != 1 or value_cache.stride(3) != 1:
        raise RuntimeError("reshape_and_cache_flash: cache head dimension must be contiguous")
    if key_cache.stride(0) != value_cache.stride(0):
        raise RuntimeError("key_cache.stride(0) == value_cache.stride(0) required")
    if key_cache.stride()[:3] != value_cache.stride()[:3]:
        raise RuntimeError("reshape_and_cache_flash: key and value cache strides must match")
    if k_scale.shape != v_scale.shape:
        raise RuntimeError("k_scale and v_scale must have the same shape")
    num_heads = key.size(1)
    if k_scale.numel() not in (1, num_heads):
        raise RuntimeError("k_scale and v_scale must be of shape [1] or [num_heads]")
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous():
        raise RuntimeError("slot_mapping must be a contiguous int64 tensor")
    kv = kv_dtype_code(kv_cache_dtype)
    rc = _lib.load().b200_reshape_and_cache_flash(
        key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), slot_mapping.data_ptr(),
        slot_mapping.size(0), num_heads, key.size(2), key_cache.size(1), key.stride(0), value.stride(0),
        key_cache.stride(0), key_cache.stride(1), key_cache.stride(2), _dt(key), kv,
        k_scale.data_ptr(), v_scale.data_ptr(), 1 if k_scale.numel() > 1 else 0, _stream(key))
    _lib.check(rc, "reshape_and_cache_flash")


def compute_slot_mapping(
    query_start_loc: torch.Tensor,
    positions: torch.Tensor,
    block_table: torch.Tensor,
    block_size: int,
    slot_mapping: torch.Tensor,
    num_reqs: int | None = None,
) -> None:
    """slot[t] = block_table[req, pos // bs] * bs + pos % bs; tail of slot_mapping = -1
    (vllm/v1/worker/block_table.py:201-229,412-475)."""
    _require_cuda(query_start_loc, positions, block_table, slot_mapping)
    if num_reqs is None:
        num_reqs = query_start_loc.numel() - 1
    if query_start_loc.dtype != torch.int32 or positions.dtype != torch.int64 or block_table.dtype != torch.int32:
        raise RuntimeError("compute_slot_mapping: query_start_loc/block_table int32 and positions int64 expected")
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous():
        raise RuntimeError("slot_mapping must be a contiguous int64 tensor")
    rc = _lib.load().b200_compute_slot_mapping(
        slot_mapping.data_ptr(), slot_mapping.numel(), positions.numel(), query_start_loc.data_ptr(), num_reqs,
        positions.data_ptr(), block_table.data_ptr(), block_table.stride(0), block_size, _stream(slot_mapping))
    _lib.check(rc, "compute_slot_mapping")


# ----------------------------------------------------------------- paged attention decode
def paged_decode_num_splits(num_seqs: int, num_q_heads: int, num_kv_heads: int, max_seq_len: int,
                            device: torch.device | None = None, head_size: int = 128,
                            kv_cache_dtype: str = "auto") -> int:
    group = num_q_heads // num_kv_heads
    qgroups = (group + 15) // 16
    sm = torch.cuda.get_device_properties(device).multi_processor_count if torch.cuda.is_available() else 148
    return int(_lib.load().b200_paged_decode_pick_splits_ex(num_seqs, num_kv_heads * qgroups, max(1, max_seq_len), sm,
                                                            head_size, kv_dtype_code(kv_cache_dtype)))


def paged_decode_workspace_bytes(num_seqs: int, num_q_heads: int, head_size: int, num_splits: int) -> int:
    return int(_lib.load().b200_paged_decode_workspace_bytes(num_seqs, num_q_heads, head_size, num_splits))


def paged_attention_decode(
    out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    block_table: torch.Tensor,
    seq_lens: torch.Tensor,
    scale: float,
    max_seq_len: int,
    kv_cache_dtype: str = "auto",
    k_scale: torch.Tensor | None = None,
    v_scale: torch.Tensor | None = None,
    num_splits: int = 0,
    workspace: torch.Tensor | None = None,
    out_lse: torch.Tensor | None = None,
    logits_soft_cap: float = 0.0,
    sliding_window: int = 0,
    token_to_seq: torch.Tensor | None = None,
    output_scale: torch.Tensor | None = None,
) -> torch.Tensor:
    """One query token per sequence (or, with token_to_seq int32 [S], one per query token of a
    mixed batch: row s reads block_table[token_to_seq[s]] with causal length seq_lens[s]).  query/out [S, Hq, D]; caches are 4-D views
    (num_blocks, block_size, Hkv, D) with any strides; block_table int32 [S, max_blocks];
    seq_lens int32 [S].  fp8 caches are uint8 / float8 storage + per-tensor k/v scales.
    output_scale (fp32 [1], optional): fused static FP8 output quant - `out` must then be a float8_e4m3fn / uint8
    tensor of query's shape and receives e4m3(o / output_scale) (AttentionImpl.forward(output_scale=...))."""
    _require_cuda(out, query, key_cache, value_cache, block_table, seq_lens, output_scale)
    if output_scale is not None:
        if out.element_size() != 1 or output_scale.numel() != 1 or output_scale.dtype != torch.float32:
            raise RuntimeError("paged_attention_decode: output_scale needs a 1-byte (float8_e4m3fn) out tensor and a "
                               "float32 scale with one element")
    elif out.dtype != query.dtype:
        raise RuntimeError("paged_attention_decode: out must have the query dtype (or pass output_scale for fp8 output)")
    if query.dim() != 3 or out.shape != query.shape:
        raise RuntimeError("paged_attention_decode: query/out must be [num_seqs, num_heads, head_size]")
    if key_cache.dim() != 4 or value_cache.dim() != 4:
        raise RuntimeError("paged_attention_decode: caches must be 4-D (num_blocks, block_size, num_kv_heads, head_size)")
    if query.stride(2) != 1 or out.stride(2) != 1 or key_cache.stride(3) != 1 or value_cache.stride(3) != 1:
        raise RuntimeError("paged_attention_decode: innermost dimensions must be contiguous")
    if block_table.dtype != torch.int32 or seq_lens.dtype != torch.int32 or block_table.stride(1) != 1:
        raise RuntimeError("paged_attention_decode: block_table / seq_lens must be int32")
    num_seqs, num_q_heads, head_size = query.shape
    num_blocks, block_size, num_kv_heads, _ = key_cache.shape
    kv = kv_dtype_code(kv_cache_dtype)
    if kv != KV_AUTO and key_cache.element_size() != 1:
        raise RuntimeError("paged_attention_decode: fp8 kv_cache_dtype needs a 1-byte cache tensor")
    if kv == KV_AUTO and key_cache.dtype != query.dtype:
        raise RuntimeError("paged_attention_decode: kv_cache_dtype='auto' needs cache dtype == query dtype")
    for nm, sc in (("k_scale", k_scale), ("v_scale", v_scale)):
        if sc is not None and (sc.numel() != 1 or sc.dtype != torch.float32):
            # the decode kernel dequantises with ONE scale per tensor; a cache written with per-head scales
            # (reshape_and_cache_flash accepts [num_heads]) would silently be read with head 0's scale
            raise RuntimeError(f"paged_attention_decode: {nm} must be a float32 tensor with exactly one element "
                               f"(per-head KV scales are not built), got shape {tuple(sc.shape)}")
    if num_splits <= 0:
        num_splits = paged_decode_num_splits(num_seqs, num_q_heads, num_kv_heads, max_seq_len, query.device,
                                             head_size, kv_cache_dtype)
    need = paged_decode_workspace_bytes(num_seqs, num_q_heads, head_size, num_splits)
    if need and (workspace is None or workspace.numel() * workspace.element_size() < need):
        workspace = torch.empty(need, dtype=torch.uint8, device=query.device)
    p = _lib.PagedDecodeParams(
        out=out.data_ptr(), query=query.data_ptr(), key_cache=key_cache.data_ptr(),
        value_cache=value_cache.data_ptr(), block_table=block_table.data_ptr(), seq_lens=seq_lens.data_ptr(),
        k_scale=_ptr(k_scale), v_scale=_ptr(v_scale), out_lse=_ptr(out_lse), token_to_seq=_ptr(token_to_seq),
        workspace=_ptr(workspace), workspace_bytes=0 if workspace is None else workspace.numel() * workspace.element_size(),
        q_stride_seq=query.stride(0), q_stride_head=query.stride(1),
        out_stride_seq=out.stride(0), out_stride_head=out.stride(1),
        kv_block_stride=key_cache.stride(0), kv_page_stride=key_cache.stride(1), kv_head_stride=key_cache.stride(2),
        v_block_stride=value_cache.stride(0), v_page_stride=value_cache.stride(1), v_head_stride=value_cache.stride(2),
        block_table_stride=block_table.stride(0), num_blocks=num_blocks,
        num_seqs=num_seqs, num_q_heads=num_q_heads, num_kv_heads=num_kv_heads, head_size=head_size,
        block_size=block_size, max_seq_len=int(max_seq_len), q_dtype=_dt(query), kv_dtype=kv,
        num_splits=int(num_splits), scale=float(scale), logits_soft_cap=float(logits_soft_cap),
        sliding_window=int(sliding_window), output_scale=_ptr(output_scale))
    rc = _lib.load().b200_paged_attention_decode(C.byref(p), _stream(query))
    _lib.check(rc, "paged_attention_decode")
    return out


def merge_attn_states(
    output: torch.Tensor,
    prefix_output: torch.Tensor,
    prefix_lse: torch.Tensor,
    suffix_output: torch.Tensor,
    suffix_lse: torch.Tensor,
    output_lse: torch.Tensor | None = None,
) -> None:
    """outputs [T, H, D]; LSEs fp32 [H, T] (vllm/_custom_ops.py:178-198)."""
    _require_cuda(output, prefix_output, prefix_lse, suffix_output, suffix_lse)
    T, H, D = output.shape
    if prefix_lse.dtype != torch.float32 or suffix_lse.dtype != torch.float32:
        raise RuntimeError("merge_attn_states: LSEs must be float32")
    if prefix_lse.shape != (H, T) or suffix_lse.shape != (H, T) or prefix_lse.stride() != suffix_lse.stride():
        raise RuntimeError("merge_attn_states: LSEs must be [num_heads, num_tokens] with equal strides")
    if output_lse is not None and output_lse.stride() != prefix_lse.stride():
        raise RuntimeError("merge_attn_states: output_lse must share the LSE layout")
    if prefix_lse.stride(1) != 1:
        raise RuntimeError("merge_attn_states: LSE token dimension must be contiguous")
    rc = _lib.load().b200_merge_attn_states(
        output.data_ptr(), _ptr(output_lse), prefix_output.data_ptr(), prefix_lse.data_ptr(),
        suffix_output.data_ptr(), suffix_lse.data_ptr(), T, H, D, output.stride(0), output.stride(1),
        prefix_output.stride(0), prefix_output.stride(1), suffix_output.stride(0), suffix_output.stride(1),
        prefix_lse.stride(0), _dt(output), _stream(output))
    _lib.check(rc, "merge_attn_states")


# --------------------------------------------------------------------------- FP8 quant
def scaled_fp8_quant(
    input: torch.Tensor,
    scale: torch.Tensor | None = None,
    num_token_padding: int | None = None,
    scale_ub: torch.Tensor | None = None,
    use_per_token_if_dynamic: bool = False,
    output: torch.Tensor | None = None,
    group_shape: tuple[int, int] | None = None,
    scale_out: torch.Tensor | None = None,
) -> tuple[torch.Tensor, torch.Tensor]:
    """Same contract as vllm/_custom_ops.py:1832-1899: scale given -> static; else
    dynamic per-tensor, or dynamic per-token when use_per_token_if_dynamic.  `scale_out`
    (ours) lets the torch.ops `Tensor! scale` variants write a caller-allocated scale."""
    _require_cuda(input, scale, scale_ub, output)
    if input.dim() != 2:
        raise RuntimeError("scaled_fp8_quant: input must be 2-D")
    if input.stride(-1) != 1:
        raise RuntimeError("last dimension of input must be contiguous")
    M, K = input.shape
    rows = max(num_token_padding, M) if num_token_padding else M
    if output is None:
        output = torch.empty((rows, K), device=input.device, dtype=FP8_DTYPE)
    elif output.dtype != FP8_DTYPE or output.stride(-1) != 1:
        raise RuntimeError("scaled_fp8_quant: output must be float8_e4m3fn with contiguous rows")
    lib, st, dt = _lib.load(), _stream(input), _dt(input)
    if scale is None:
        if use_per_token_if_dynamic:
            scale = scale_out if scale_out is not None else torch.empty((M, 1), device=input.device, dtype=torch.float32)
            if scale.numel() != M or scale.dtype != torch.float32 or not scale.is_contiguous():
                raise RuntimeError("dynamic_per_token_scaled_fp8_quant: scale must be contiguous float32 with one entry per token")
            rc = lib.b200_dynamic_per_token_scaled_fp8_quant(
                output.data_ptr(), input.data_ptr(), scale.data_ptr(), _ptr(scale_ub), M, K, input.stride(0),
                output.stride(0), dt, st)
            _lib.check(rc, "dynamic_per_token_scaled_fp8_quant")
        else:
            scale = scale_out if scale_out is not None else torch.empty(1, device=input.device, dtype=torch.float32)
            if scale.numel() != 1 or scale.dtype != torch.float32:
                raise RuntimeError("dynamic_scaled_fp8_quant: scale must be a float32 tensor with one element")
            rc = lib.b200_dynamic_scaled_fp8_quant(output.data_ptr(), input.data_ptr(), scale.data_ptr(), M, K,
                                                   input.stride(0), output.stride(0), dt, st)
            _lib.check(rc, "dynamic_scaled_fp8_quant")
        return output, scale
    if scale.dtype != torch.float32:
        raise RuntimeError("scaled_fp8_quant: scale must be float32")
    gm, gn, si, sj = _static_scale_layout(scale, M, K, group_shape)
    rc = lib.b200_static_scaled_fp8_quant(output.data_ptr(), input.data_ptr(), scale.data_ptr(), M, K,
                                          input.stride(0), output.stride(0), gm, gn, si, sj, dt, st)
    _lib.check(rc, "static_scaled_fp8_quant")
    return output, scale


def _static_scale_layout(scale: torch.Tensor, M: int, K: int, group_shape):
    """Scale-shape rules of static_scaled_fp8_quant (fp8/common.cu:199-283)."""
    if scale.dim() == 0 or scale.numel() == 1:
        return max(M, 1), K, 0, 0
    if scale.dim() == 1:
        if group_shape is None or len(group_shape) != 2:
            raise RuntimeError("1D scale requires explicit group_shape to disambiguate per-channel vs per-token "
                               "quantization. Use group_shape=(-1, 1) for per-channel or group_shape=(1, -1) for per-token.")
        gm = M if group_shape[0] == -1 else int(group_shape[0])
        gn = K if group_shape[1] == -1 else int(group_shape[1])
        em, en = M // gm, K // gn
        if scale.numel() != em * en:
            raise RuntimeError(f"1D scale length ({scale.numel()}) does not match expected size ({em * en})")
        if em == 1:
            return gm, gn, 0, scale.stride(0)
        if en == 1:
            return gm, gn, scale.stride(0), 0
        raise RuntimeError("1D scale can only be used when one of the scale dimensions is 1. "
                           "For 2D group scaling, use a 2D scale tensor.")
    if scale.dim() == 2:
        s0, s1 = scale.shape
        if M % s0 or K % s1:
            raise RuntimeError("num_tokens / hidden_size must be divisible by the scale dimensions")
        gm, gn = M // s0, K // s1
        if group_shape is not None:
            egm = M if group_shape[0] == -1 else int(group_shape[0])
            egn = K if group_shape[1] == -1 else int(group_shape[1])
            if (egm, egn) != (gm, gn):
                raise RuntimeError("Explicit group_shape does not match inferred group shape from 2D scale tensor")
        return gm, gn, scale.stride(0), scale.stride(1)
    raise RuntimeError(f"scale must be 0D, 1D, or 2D tensor, but got {scale.dim()}D")


# --------------------------------------------------------------------------- FP8 GEMM
_mm_workspaces: dict[tuple[int, int], torch.Tensor] = {}


def _mm_workspace(device: torch.device, nbytes: int) -> torch.Tensor | None:
    """Stream-K fixup scratch of the decode-regime GEMM (flags zero-initialised once; the kernels
    leave them zeroed).  One buffer per (device, stream): GEMMs on one stream are serialised, GEMMs
    on different streams must not share partials."""
    if nbytes <= 0:
        return None
    key = (device.index or 0, torch.cuda.current_stream(device).cuda_stream)
    ws = _mm_workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        if torch.cuda.is_current_stream_capturing():
            # never allocate/zero inside a capture: fall back to the capture-origin stream's buffer if any,
            # else run without stream-K (whole tiles per CTA)
            for (d, _), w in _mm_workspaces.items():
                if d == (device.index or 0) and w.numel() >= nbytes:
                    return w
            return None
        ws = torch.zeros(max(nbytes, 24 << 20), dtype=torch.uint8, device=device)
        _mm_workspaces[key] = ws
    return ws


def cutlass_scaled_mm(
    a: torch.Tensor,
    b: torch.Tensor,
    scale_a: torch.Tensor,
    scale_b: torch.Tensor,
    out_dtype: torch.dtype,
    bias: torch.Tensor | None = None,
    out: torch.Tensor | None = None,
) -> torch.Tensor:
    """out = (scale_a * a) @ (scale_b * b) + bias; a [M,K] row-major e4m3, b [K,N]
    column-major (b.stride(0) == 1), scales fp32 with 1 | M and 1 | N elements
    (vllm/_custom_ops.py:756-804; checks scaled_mm_entry.cu:203-216)."""
    _require_cuda(a, b, scale_a, scale_b, bias)
    if a.dim() != 2 or b.dim() != 2 or a.shape[1] != b.shape[0]:
        raise RuntimeError("cutlass_scaled_mm: a [M,K] and b [K,N] expected")
    if out_dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("cutlass_scaled_mm: out_dtype must be bfloat16 or float16")
    if a.dtype != FP8_DTYPE or b.dtype != FP8_DTYPE:
        raise RuntimeError("cutlass_scaled_mm: this build implements float8_e4m3fn operands")
    M, K = a.shape
    N = b.shape[1]
    if b.shape[0] % 16 or b.shape[1] % 16:
        raise RuntimeError("cutlass_scaled_mm: b.shape[0] and b.shape[1] must be multiples of 16")
    if a.stride(1) != 1 or b.stride(0) != 1:
        raise RuntimeError("cutlass_scaled_mm: a must be row-major and b column-major")
    if scale_a.dtype != torch.float32 or scale_b.dtype != torch.float32:
        raise RuntimeError("cutlass_scaled_mm: scales must be float32")
    if scale_a.numel() not in (1, M) or scale_b.numel() not in (1, N):
        raise RuntimeError("cutlass_scaled_mm: scale_a needs 1 or M elements and scale_b 1 or N")
    if not scale_a.is_contiguous() or not scale_b.is_contiguous():
        raise RuntimeError("cutlass_scaled_mm: scales must be contiguous")
    if bias is not None and (bias.numel() != N or bias.dtype != out_dtype or not bias.is_contiguous()):
        raise RuntimeError("cutlass_scaled_mm: bias must be a contiguous [N] tensor of out_dtype")
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    elif out.shape != (M, N) or out.dtype != out_dtype or out.stride(1) != 1:
        raise RuntimeError("cutlass_scaled_mm: out must be a row-major [M, N] tensor of out_dtype")
    lib = _lib.load()
    ws = _mm_workspace(a.device, int(lib.b200_scaled_mm_workspace_bytes(M, N)))
    rc = lib.b200_scaled_mm_fp8(
        out.data_ptr(), a.data_ptr(), b.data_ptr(), scale_a.data_ptr(), scale_b.data_ptr(), _ptr(bias), M, N, K,
        a.stride(0), b.stride(1), out.stride(0), scale_a.numel(), scale_b.numel(), _DT[out_dtype],
        _ptr(ws), 0 if ws is None else ws.numel(), _stream(a))
    _lib.check(rc, "cutlass_scaled_mm")
    return out


# ------------------------------------------------------------------------- NVFP4 quant
def round_up(x: int, y: int) -> int:
    return (x + y - 1) // y * y


def create_fp4_scale_tensor(m: int, n: int, device, is_sf_swizzled_layout: bool) -> torch.Tensor:
    """vllm/_custom_ops.py:36-70 (swizzled buffer: int32 [ru(m,128), ru(n/16,4)/4])."""
    if is_sf_swizzled_layout:
        return torch.zeros((round_up(m, 128), round_up(n // 16, 4) // 4), device=device, dtype=torch.int32)
    return torch.empty((m, n // 16), device=device, dtype=torch.uint8)


def scaled_fp4_quant(
    input: torch.Tensor,
    input_global_scale: torch.Tensor,
    is_sf_swizzled_layout: bool = True,
    padded_n: int | None = None,
) -> tuple[torch.Tensor, torch.Tensor]:
    """Returns (uint8 [m, n/2] packed e2m1, float8_e4m3fn scale factors), same as
    vllm/_custom_ops.py:1523-1598 (128x4 swizzled SF layout by default)."""
    _require_cuda(input, input_global_scale)
    if input.dim() < 1:
        raise RuntimeError(f"input.ndim needs to be >= 1, but got {input.dim()}.")
    input = input.reshape(1 if input.dim() == 1 else -1, input.shape[-1])
    m, n = input.shape
    if n % 16:
        raise RuntimeError(f"last dim has to be multiple of 16, but got {n}.")
    if input.dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError(f"input.dtype needs to be fp16 or bf16 but got {input.dtype}.")
    out_n = n if padded_n is None else padded_n
    if out_n < n or out_n % 16:
        raise RuntimeError("padded_n must be >= n and a multiple of 16")
    if not input.is_contiguous():
        input = input.contiguous()
    output = torch.empty((m, out_n // 2), device=input.device, dtype=torch.uint8)
    output_scale = create_fp4_scale_tensor(m, out_n, input.device, is_sf_swizzled_layout)
    rc = _lib.load().b200_scaled_fp4_quant(
        output.data_ptr(), output_scale.data_ptr(), input.data_ptr(), input_global_scale.data_ptr(), m, n, out_n,
        1 if is_sf_swizzled_layout else 0, _dt(input), _stream(input))
    _lib.check(rc, "scaled_fp4_quant")
    return output, output_scale.view(torch.float8_e4m3fn)


def scaled_fp4_quant_out(input: torch.Tensor, input_global_scale: torch.Tensor, is_sf_swizzled_layout: bool,
                         output: torch.Tensor, output_scale: torch.Tensor) -> None:
    """The `.out` overload (torch_bindings.cpp:218-222): caller-allocated packed output [m, out_n/2]
    and scale buffer (swizzled buffers must be zero-initialised by the caller or are fully
    rewritten here: every padded SF is stored)."""
    _require_cuda(input, input_global_scale, output, output_scale)
    input = input.reshape(1 if input.dim() == 1 else -1, input.shape[-1])
    m, n = input.shape
    if n % 16:
        raise RuntimeError(f"last dim has to be multiple of 16, but got {n}.")
    if input.dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError(f"input.dtype needs to be fp16 or bf16 but got {input.dtype}.")
    if output.dtype != torch.uint8 or output.dim() != 2 or output.shape[0] != m or not output.is_contiguous():
        raise RuntimeError("scaled_fp4_quant.out: output must be contiguous uint8 [m, out_n/2]")
    out_n = output.shape[1] * 2
    if out_n < n or out_n % 16:
        raise RuntimeError("scaled_fp4_quant.out: output width must cover n and be a multiple of 16 elements")
    need = (round_up(m, 128) * round_up(out_n // 16, 4)) if is_sf_swizzled_layout else m * (out_n // 16)
    if output_scale.numel() * output_scale.element_size() < need or not output_scale.is_contiguous():
        raise RuntimeError("scaled_fp4_quant.out: output_scale buffer too small or not contiguous")
    if not input.is_contiguous():
        input = input.contiguous()
    rc = _lib.load().b200_scaled_fp4_quant(
        output.data_ptr(), output_scale.data_ptr(), input.data_ptr(), input_global_scale.data_ptr(), m, n, out_n,
        1 if is_sf_swizzled_layout else 0, _dt(input), _stream(input))
    _lib.check(rc, "scaled_fp4_quant")


def silu_and_mul_nvfp4_quant(result: torch.Tensor, result_block_scale: torch.Tensor, input: torch.Tensor,
                             input_global_scale: torch.Tensor) -> None:
    """torch.ops._C.silu_and_mul_nvfp4_quant (torch_bindings.cpp:251-253): input [m, 2d] = [gate | up];
    result uint8 [m, d/2]; result_block_scale = the swizzled scale buffer for [m, d] (fully rewritten)."""
    _require_cuda(input, input_global_scale, result, result_block_scale)
    input = input.reshape(-1, input.shape[-1])
    m, d2 = input.shape
    if d2 % 32:
        raise RuntimeError(f"The N dimension must be multiple of 16. (got 2d = {d2})")
    d = d2 // 2
    if input.dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError(f"Unsupported input data type for quantize_to_fp4: {input.dtype}")
    if result.dtype != torch.uint8 or result.numel() != m * d // 2 or not result.is_contiguous():
        raise RuntimeError("silu_and_mul_nvfp4_quant: result must be contiguous uint8 [m, d/2]")
    need = round_up(m, 128) * round_up(d // 16, 4)
    if result_block_scale.numel() * result_block_scale.element_size() < need or not result_block_scale.is_contiguous():
        raise RuntimeError("silu_and_mul_nvfp4_quant: result_block_scale buffer too small or not contiguous")
    if not input.is_contiguous():
        input = input.contiguous()
    rc = _lib.load().b200_silu_and_mul_nvfp4_quant(result.data_ptr(), result_block_scale.data_ptr(), input.data_ptr(),
                                                   input_global_scale.data_ptr(), m, d, _dt(input), _stream(input))
    _lib.check(rc, "silu_and_mul_nvfp4_quant")


def rms_norm_nvfp4_quant(input: torch.Tensor, weight: torch.Tensor, epsilon: float, input_global_scale: torch.Tensor,
                         residual: torch.Tensor | None = None, result: torch.Tensor | None = None,
                         result_block_scale: torch.Tensor | None = None) -> tuple[torch.Tensor, torch.Tensor]:
    """(residual += input in place;) NVFP4 quant of RMSNorm(.) * weight with swizzled scales - one kernel for
    fused_add_rms_norm + scaled_fp4_quant (no reference counterpart; bit-identical to that composition of ours).
    Returns (uint8 [m, h/2], float8_e4m3fn scales)."""
    _require_cuda(input, weight, input_global_scale)
    h = input.shape[-1]
    x = input.view(-1, h) if input.is_contiguous() else input.reshape(-1, h)
    m = x.shape[0]
    if h % 16:
        raise RuntimeError(f"last dim has to be multiple of 16, but got {h}.")
    if x.dtype not in (torch.float16, torch.bfloat16) or weight.dtype != x.dtype:
        raise RuntimeError("rms_norm_nvfp4_quant: input and weight must both be fp16 or bf16")
    if x.stride(-1) != 1:
        x = x.contiguous()
    if residual is not None and (residual.dtype != x.dtype or residual.numel() != m * h or not residual.is_contiguous()):
        raise RuntimeError("rms_norm_nvfp4_quant: residual must be contiguous, same dtype and shape as input")
    if result is None:
        result = torch.empty((m, h // 2), device=x.device, dtype=torch.uint8)
    if result_block_scale is None:
        result_block_scale = create_fp4_scale_tensor(m, h, x.device, True)
    rc = _lib.load().b200_rms_norm_nvfp4_quant(
        result.data_ptr(), result_block_scale.data_ptr(), x.data_ptr(), weight.data_ptr(),
        residual.data_ptr() if residual is not None else None, input_global_scale.data_ptr(), m, h, x.stride(0),
        float(epsilon), _dt(x), _stream(x))
    _lib.check(rc, "rms_norm_nvfp4_quant")
    return result, result_block_scale.view(torch.float8_e4m3fn)


def cutlass_scaled_fp4_mm(
    a: torch.Tensor,
    b: torch.Tensor,
    block_scale_a: torch.Tensor,
    block_scale_b: torch.Tensor,
    alpha: torch.Tensor,
    out_dtype: torch.dtype,
    out: torch.Tensor | None = None,
) -> torch.Tensor:
    """a [M, K/2] u8, b [N, K/2] u8, swizzled e4m3 block scales, alpha fp32 scalar
    (vllm/_custom_ops.py:733-745; checks nvfp4_scaled_mm_kernels.cu:279-347)."""
    _require_cuda(a, b, block_scale_a, block_scale_b, alpha)
    if a.dim() != 2 or b.dim() != 2:
        raise RuntimeError("a and b must be matrices")
    if a.shape[1] != b.shape[1]:
        raise RuntimeError(f"a and b shapes cannot be multiplied ({tuple(a.shape)} and {tuple(b.shape)})")
    if a.dtype != torch.uint8 or b.dtype != torch.uint8:
        raise RuntimeError("a and b must be packed uint8 (two e2m1 per byte)")
    if not a.is_contiguous() or not b.is_contiguous():
        raise RuntimeError("a and b must be contiguous")
    m, n, k = a.shape[0], b.shape[0], a.shape[1] * 2
    if k % 32:
        raise RuntimeError(f"Expected k to be divisible by 32, but got a shape: {tuple(a.shape)}")
    if n % 32:
        raise RuntimeError(f"Expected n to be divisible by 32, but got b shape: {tuple(b.shape)}")
    sf_k = round_up(k // 16, 4)
    sa = block_scale_a.view(torch.uint8) if block_scale_a.element_size() == 1 else block_scale_a.view(torch.uint8)
    sb = block_scale_b.view(torch.uint8) if block_scale_b.element_size() == 1 else block_scale_b.view(torch.uint8)
    if sa.numel() != round_up(m, 128) * sf_k:
        raise RuntimeError(f"scale_a must be padded and swizzled to a shape ({round_up(m, 128)}, {sf_k}), "
                           f"but got {sa.numel()} elements")
    if sb.numel() != round_up(n, 128) * sf_k:
        raise RuntimeError(f"scale_b must be padded and swizzled to a shape ({round_up(n, 128)}, {sf_k}), "
                           f"but got {sb.numel()} elements")
    if not sa.is_contiguous() or not sb.is_contiguous():
        raise RuntimeError("block scales must be contiguous")
    if alpha.dtype != torch.float32:
        raise RuntimeError("alpha must be float32")
    if out_dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("cutlass_scaled_fp4_mm: out_dtype must be bfloat16 or float16")
    if out is None:
        out = torch.empty((m, n), dtype=out_dtype, device=a.device)
    elif out.shape != (m, n) or out.dtype != out_dtype or out.stride(1) != 1:
        raise RuntimeError("cutlass_scaled_fp4_mm: out must be a row-major [M, N] tensor of out_dtype")
    rc = _lib.load().b200_scaled_mm_nvfp4(out.data_ptr(), a.data_ptr(), b.data_ptr(), sa.data_ptr(), sb.data_ptr(),
                                          alpha.data_ptr(), m, n, k, out.stride(0), _DT[out_dtype], _stream(a))
    _lib.check(rc, "cutlass_scaled_fp4_mm")
    return out


# ------------------------------------------------------------------ element-wise companions
def rms_norm(out: torch.Tensor, input: torch.Tensor, weight: torch.Tensor, epsilon: float) -> None:
    _require_cuda(out, input, weight)
    x = input.reshape(-1, input.shape[-1])
    if x.stride(-1) != 1:
        raise RuntimeError("rms_norm: input rows must be contiguous")
    if out.dtype != input.dtype or out.numel() != input.numel() or not out.is_contiguous():
        raise RuntimeError("rms_norm: out must be a contiguous tensor of the input's dtype and size")
    rc = _lib.load().b200_rms_norm(out.data_ptr(), x.data_ptr(), weight.data_ptr(), None, x.shape[0], x.shape[1],
                                   x.stride(0), float(epsilon), _dt(x), _stream(x))
    _lib.check(rc, "rms_norm")


def fused_add_rms_norm(input: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor, epsilon: float) -> None:
    """In place like the reference: residual += input (rounded), input = norm(residual) * w."""
    _require_cuda(input, residual, weight)
    # in place: a reshape() of a non-contiguous input would silently normalise a temporary copy
    if not input.is_contiguous() or not residual.is_contiguous():
        raise RuntimeError("fused_add_rms_norm: contiguous tensors expected")
    x = input.view(-1, input.shape[-1])
    rc = _lib.load().b200_rms_norm(x.data_ptr(), x.data_ptr(), weight.data_ptr(), residual.data_ptr(), x.shape[0],
                                   x.shape[1], x.stride(0), float(epsilon), _dt(x), _stream(x))
    _lib.check(rc, "fused_add_rms_norm")


def silu_and_mul(out: torch.Tensor, input: torch.Tensor) -> None:
    _require_cuda(out, input)
    x = input.reshape(-1, input.shape[-1])
    if not x.is_contiguous() or not out.is_contiguous():
        raise RuntimeError("silu_and_mul: contiguous tensors expected")
    rc = _lib.load().b200_silu_and_mul(out.data_ptr(), x.data_ptr(), x.shape[0], x.shape[1] // 2, _dt(x), _stream(x))
    _lib.check(rc, "silu_and_mul")


def rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                     cos_sin_cache: torch.Tensor, is_neox: bool = True) -> None:
    _require_cuda(positions, query, key, cos_sin_cache)
    if not is_neox:
        raise RuntimeError("rotary_embedding: only the neox layout is built")
    T = positions.numel()
    # in place: view() raises where reshape() would hand back a copy (rows may be strided slices of a fused qkv)
    q = query.view(T, -1)
    k = key.view(T, -1)
    if q.stride(1) != 1 or k.stride(1) != 1:
        raise RuntimeError("rotary_embedding: query/key rows must be contiguous")
    rc = _lib.load().b200_rotary_embedding_neox(
        q.data_ptr(), k.data_ptr(), positions.data_ptr(), cos_sin_cache.data_ptr(), T, q.shape[1] // head_size,
        k.shape[1] // head_size, head_size, cos_sin_cache.shape[1], q.stride(0), k.stride(0), _dt(q), _stream(q))
    _lib.check(rc, "rotary_embedding")


def rms_norm_dynamic_per_token_quant(input: torch.Tensor, weight: torch.Tensor, epsilon: float,
                                     scale_ub: torch.Tensor | None = None, residual: torch.Tensor | None = None,
                                     output: torch.Tensor | None = None, scales: torch.Tensor | None = None,
                                     ) -> tuple[torch.Tensor, torch.Tensor]:
    """(fused add +) RMSNorm + FP8 per-token quant, vllm/_custom_ops.py rms_norm_dynamic_per_token_quant
    (schema torch_bindings.cpp:394-396).  residual (if given) is updated in place.  Returns (e4m3 [M, H], scales [M, 1])."""
    _require_cuda(input, weight, scale_ub, residual, output, scales)
    x = input.reshape(-1, input.shape[-1])
    if x.stride(-1) != 1:
        raise RuntimeError("Input must be contiguous in the last dimension")
    if weight.dtype != x.dtype or (residual is not None and (residual.dtype != x.dtype or not residual.is_contiguous())):
        raise RuntimeError("rms_norm_dynamic_per_token_quant: weight/residual dtype must match input; residual contiguous")
    M, H = x.shape
    if output is None:
        output = torch.empty((M, H), device=x.device, dtype=FP8_DTYPE)
    elif output.dtype != FP8_DTYPE or not output.is_contiguous():
        raise RuntimeError("rms_norm_dynamic_per_token_quant: output must be contiguous float8_e4m3fn")
    if scales is None:
        scales = torch.empty((M, 1), device=x.device, dtype=torch.float32)
    rc = _lib.load().b200_rms_norm_dynamic_per_token_quant(
        output.data_ptr(), x.data_ptr(), weight.data_ptr(), scales.data_ptr(), float(epsilon), _ptr(scale_ub),
        _ptr(residual), M, H, x.stride(0), _dt(x), _stream(x))
    _lib.check(rc, "rms_norm_dynamic_per_token_quant")
    return output, scales


def silu_and_mul_dynamic_per_token_quant(input: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    """SiLU(x[:, :d]) * x[:, d:] then dynamic per-token FP8 quant in one launch (bit-identical to
    silu_and_mul + scaled_fp8_quant(use_per_token_if_dynamic=True))."""
    _require_cuda(input)
    x = input.reshape(-1, input.shape[-1])
    if not x.is_contiguous():
        raise RuntimeError("silu_and_mul_dynamic_per_token_quant: contiguous input expected")
    M, d = x.shape[0], x.shape[1] // 2
    out = torch.empty((M, d), device=x.device, dtype=FP8_DTYPE)
    scales = torch.empty((M, 1), device=x.device, dtype=torch.float32)
    rc = _lib.load().b200_silu_and_mul_dynamic_per_token_quant(out.data_ptr(), scales.data_ptr(), x.data_ptr(), M, d,
                                                               _dt(x), _stream(x))
    _lib.check(rc, "silu_and_mul_dynamic_per_token_quant")
    return out, scales


def rotary_embedding_and_cache_write(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                                     value: torch.Tensor, head_size: int, cos_sin_cache: torch.Tensor,
                                     key_cache: torch.Tensor, value_cache: torch.Tensor, slot_mapping: torch.Tensor,
                                     kv_cache_dtype: str, k_scale: torch.Tensor, v_scale: torch.Tensor,
                                     is_neox: bool = True) -> None:
    """RoPE on query/key in place + KV-cache write in one launch (the fused op behind
    AttentionImpl.do_rope_and_kv_cache_update, vllm/v1/attention/backend.py:932-980)."""
    _require_cuda(positions, query, key, value, cos_sin_cache, key_cache, value_cache, slot_mapping)
    if not is_neox:
        raise RuntimeError("rotary_embedding_and_cache_write: only the neox layout is built")
    T = positions.numel()
    q, k, v = query.view(T, -1), key.view(T, -1), value.view(T, -1)  # in place: never a reshape() copy
    if q.stride(1) != 1 or k.stride(1) != 1 or v.stride(1) != 1 or key_cache.stride(3) != 1:
        raise RuntimeError("rotary_embedding_and_cache_write: innermost dimensions must be contiguous")
    if key_cache.stride() != value_cache.stride():
        raise RuntimeError("rotary_embedding_and_cache_write: key and value cache strides must match")
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous():
        raise RuntimeError("slot_mapping must be a contiguous int64 tensor")
    hkv = k.shape[1] // head_size
    if k_scale.numel() not in (1, hkv) or k_scale.shape != v_scale.shape:
        raise RuntimeError("k_scale and v_scale must be of shape [1] or [num_heads]")
    rc = _lib.load().b200_rotary_embedding_and_cache_write(
        q.data_ptr(), k.data_ptr(), v.data_ptr(), positions.data_ptr(), cos_sin_cache.data_ptr(), key_cache.data_ptr(),
        value_cache.data_ptr(), slot_mapping.data_ptr(), T, slot_mapping.numel(), q.shape[1] // head_size, hkv,
        head_size, cos_sin_cache.shape[1], q.stride(0), k.stride(0), v.stride(0), key_cache.size(1),
        key_cache.stride(0), key_cache.stride(1), key_cache.stride(2), _dt(q), kv_dtype_code(kv_cache_dtype),
        k_scale.data_ptr(), v_scale.data_ptr(), 1 if k_scale.numel() > 1 else 0, _stream(q))
    _lib.check(rc, "rotary_embedding_and_cache_write")


def swizzle_blockscale(scale: torch.Tensor) -> torch.Tensor:
    """Linear [M, K/16] e4m3 block scales -> padded 128x4-swizzled layout
    (vllm/model_executor/layers/quantization/utils/nvfp4_utils.py:13-53). Pure index shuffle."""
    assert scale.dtype == torch.float8_e4m3fn and scale.dim() == 2
    M, K = scale.shape
    Mp, Kp = round_up(M, 128), round_up(K, 4)
    padded = torch.zeros((Mp, Kp), dtype=torch.uint8, device=scale.device)
    padded[:M, :K] = scale.view(torch.uint8)
    sw = padded.reshape(Mp // 128, 4, 32, Kp // 4, 4).permute(0, 3, 2, 1, 4).contiguous()
    return sw.reshape(Mp, Kp).view(torch.float8_e4m3fn)
