"""CPU ORACLE — TEST INFRASTRUCTURE ONLY.

A plain torch/numpy restatement of the reference's algorithms for the decode hot
path.  Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline /
`--impl reference` legs may import this module, and only as the *checker*; the
product package `vllm_b200` never imports it and has no CPU fallback.

Pinning: every function below is checked in `tests/test_oracle_golden.py` against
fixtures in `tests/golden/` that were produced by running the reference's own
Python test references (imported from /root/reference) and the reference's CPU
attention kernels compiled unmodified into `oracle/_ref/` — see
`oracle/make_golden.py` (the generating script) and `oracle/build_ref.py`.

Each function cites the reference code (path:line under /root/reference) it follows.
"""
from __future__ import annotations

import math

import torch

FP8_E4M3_MAX = 448.0
FP8_E5M2_MAX = 57344.0
PAD_SLOT_ID = -1  # vllm/v1/attention/backends/utils.py:43


# --------------------------------------------------------------------------- helpers
def to_fp8(x: torch.Tensor, kind: str = "e4m3") -> torch.Tensor:
    """fp32 -> fp8 with round-to-nearest-even and saturation to the finite range
    (== __nv_cvt_float_to_fp8(x, __NV_SATFINITE, kind),
    csrc/quantization/w8a8/fp8/nvidia/quant_utils.cuh:480-503).  NaN stays NaN."""
    if kind == "e4m3":
        return torch.clamp(x.float(), -FP8_E4M3_MAX, FP8_E4M3_MAX).to(torch.float8_e4m3fn)
    return torch.clamp(x.float(), -FP8_E5M2_MAX, FP8_E5M2_MAX).to(torch.float8_e5m2)


def clamp_to_e4m3(x: torch.Tensor) -> torch.Tensor:
    """scaled_fp8_conversion (csrc/quantization/w8a8/fp8/common.cuh:57-76):
    r = fmaxf(-448, fminf(x, 448)) then the hardware cvt.  fminf/fmaxf drop a NaN operand,
    so NaN (0 * inf when a whole tensor is zero) becomes +448 — unlike torch.clamp."""
    hi = torch.tensor(FP8_E4M3_MAX, dtype=torch.float32)
    r = torch.fmax(-hi, torch.fmin(x.float(), hi))
    return r.to(torch.float8_e4m3fn)


def fp8_kind(kv_cache_dtype: str) -> str | None:
    """csrc/attention/dtype_fp8.cuh:21 get_fp8_kv_cache_data_type."""
    if kv_cache_dtype in ("auto", "bfloat16", "float16", "half"):
        return None
    if kv_cache_dtype in ("fp8", "fp8_e4m3"):
        return "e4m3"
    if kv_cache_dtype == "fp8_e5m2":
        return "e5m2"
    raise ValueError(f"Unsupported data type of kv cache: {kv_cache_dtype}")


# ----------------------------------------------------------------------- slot mapping
def compute_slot_mapping(query_start_loc, positions, block_table, block_size, max_num_tokens):
    """vllm/v1/worker/block_table.py:412-475 (no context parallelism):
    slot = block_table[req, pos // bs] * bs + pos % bs; slots [num_tokens, max) = -1."""
    num_reqs = query_start_loc.numel() - 1
    num_tokens = positions.numel()
    slots = torch.empty(max_num_tokens, dtype=torch.int64)
    slots[num_tokens:] = PAD_SLOT_ID
    for r in range(num_reqs):
        s, e = int(query_start_loc[r]), int(query_start_loc[r + 1])
        pos = positions[s:e].to(torch.int64)
        blk = block_table[r, pos // block_size].to(torch.int64)
        slots[s:e] = blk * block_size + pos % block_size
    return slots


# ------------------------------------------------------------------------- cache write
def reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale):
    """csrc/libtorch_stable/cache_kernels.cu:320-406 (kernel), :801-871 (host).
    In place on key_cache/value_cache, 4-D views (num_blocks, block_size, H, D) of any
    stride; fp8 caches are uint8 storage.  Token count = slot_mapping.size(0)."""
    kind = fp8_kind(kv_cache_dtype)
    block_size = key_cache.size(1)
    H = key.size(1)
    ks = k_scale.float().reshape(-1)
    vs = v_scale.float().reshape(-1)
    ks = ks.expand(H) if ks.numel() == 1 else ks
    vs = vs.expand(H) if vs.numel() == 1 else vs
    for t in range(slot_mapping.size(0)):
        slot = int(slot_mapping[t])
        if slot < 0:  # cache_kernels.cu:333-337
            continue
        b, o = slot // block_size, slot % block_size
        if kind is None:
            key_cache[b, o] = key[t]
            value_cache[b, o] = value[t]
        else:
            kq = to_fp8(key[t].float() / ks[:, None], kind).view(torch.uint8)
            vq = to_fp8(value[t].float() / vs[:, None], kind).view(torch.uint8)
            key_cache[b, o] = kq if key_cache.dtype == torch.uint8 else kq.view(key_cache.dtype)
            value_cache[b, o] = vq if value_cache.dtype == torch.uint8 else vq.view(value_cache.dtype)


def dequant_cache(cache: torch.Tensor, kv_cache_dtype: str, scale) -> torch.Tensor:
    kind = fp8_kind(kv_cache_dtype)
    if kind is None:
        return cache.float()
    dt = torch.float8_e4m3fn if kind == "e4m3" else torch.float8_e5m2
    c = cache.view(dt) if cache.dtype == torch.uint8 else cache
    return c.float() * float(scale)


# -------------------------------------------------------------------- paged attention
def paged_attention_decode(query, key_cache, value_cache, block_table, seq_lens, scale, kv_cache_dtype="auto",
                           k_scale=1.0, v_scale=1.0, soft_cap=0.0, sliding_window=0, return_lse=False):
    """Single-query attention over a paged cache, fp32 math.  Follows
    tests/kernels/attention/test_triton_unified_attention.py:35-91 (ref_paged_attn) and
    tests/kernels/attention/test_attention.py:44-111 (ref_single_query_cached_kv_attention).
    query [S, Hq, D]; caches (num_blocks, block_size, Hkv, D)."""
    S, Hq, D = query.shape
    _, bs, Hkv, _ = key_cache.shape
    G = Hq // Hkv
    out = torch.zeros((S, Hq, D), dtype=torch.float32)
    lse = torch.full((S, Hq), -math.inf, dtype=torch.float32)
    for s in range(S):
        L = int(seq_lens[s])
        if L == 0:
            continue
        nb = (L + bs - 1) // bs
        blocks = block_table[s, :nb].long()
        k = dequant_cache(key_cache[blocks], kv_cache_dtype, k_scale).reshape(nb * bs, Hkv, D)[:L]
        v = dequant_cache(value_cache[blocks], kv_cache_dtype, v_scale).reshape(nb * bs, Hkv, D)[:L]
        lo = max(0, L - sliding_window) if sliding_window > 0 else 0
        k, v = k[lo:], v[lo:]
        k = k.repeat_interleave(G, dim=1)
        v = v.repeat_interleave(G, dim=1)
        logits = torch.einsum("hd,lhd->hl", query[s].float(), k) * scale
        if soft_cap > 0:
            logits = soft_cap * torch.tanh(logits / soft_cap)
        lse[s] = torch.logsumexp(logits, dim=-1)
        p = torch.softmax(logits, dim=-1)
        out[s] = torch.einsum("hl,lhd->hd", p, v)
    return (out, lse) if return_lse else out


def combine_splits(part_o, part_m, part_l):
    """reduce_segments, vllm/v1/attention/ops/triton_unified_attention.py:735-759:
    m = max m_i; o = sum o_i e^{m_i-m} / sum l_i e^{m_i-m} (0 where the sum is 0).
    part_o [..., S, D] un-normalised, part_m/part_l [..., S] (natural-log domain)."""
    m = part_m.max(dim=-1, keepdim=True).values
    w = torch.where(torch.isinf(part_m) & (part_m < 0), torch.zeros_like(part_m), torch.exp(part_m - m))
    l = (part_l * w).sum(-1)
    o = (part_o * w[..., None]).sum(-2)
    return torch.where(l[..., None] > 0, o / l[..., None], torch.zeros_like(o))


def merge_attn_states(prefix_output, prefix_lse, suffix_output, suffix_lse):
    """csrc/libtorch_stable/attention/merge_attn_states.cu:19-257; outputs [T,H,D],
    LSEs [H,T]; +-inf LSE -> -inf; both -inf -> prefix output and -inf."""
    pl = torch.where(torch.isinf(prefix_lse), torch.full_like(prefix_lse, -math.inf), prefix_lse)
    sl = torch.where(torch.isinf(suffix_lse), torch.full_like(suffix_lse, -math.inf), suffix_lse)
    mx = torch.maximum(pl, sl)
    dead = torch.isinf(mx)
    mx_safe = torch.where(dead, torch.zeros_like(mx), mx)
    pe, se = torch.exp(pl - mx_safe), torch.exp(sl - mx_safe)
    tot = pe + se
    pw = (pe / tot).transpose(0, 1)[..., None]
    sw = (se / tot).transpose(0, 1)[..., None]
    out = prefix_output.float() * pw + suffix_output.float() * sw
    out = torch.where(dead.transpose(0, 1)[..., None], prefix_output.float(), out)
    lse = torch.where(dead, mx, torch.log(tot) + mx_safe)
    return out.to(prefix_output.dtype), lse


# --------------------------------------------------------------------------- FP8 quant
def static_scaled_fp8_quant(x, scale):
    """fp8/common.cu:13-76 + common.cuh:57-76: e4m3(clamp(x * (1/scale)))."""
    inv = (torch.ones((), dtype=torch.float32) / scale.float())
    return clamp_to_e4m3(x.float() * inv)


def dynamic_scaled_fp8_quant(x):
    """fp8/common.cu:78-135: scale = max|x| / 448 (per tensor), out = e4m3(clamp(x * (1/scale)))."""
    scale = (x.float().abs().max() / torch.tensor(FP8_E4M3_MAX, dtype=torch.float32)).reshape(1)
    return static_scaled_fp8_quant(x, scale), scale


def dynamic_per_token_scaled_fp8_quant(x, scale_ub=None):
    """fp8/common.cu:137-179: s = max(min(amax, ub)/448, 1/(448*512)); out = e4m3(clamp(x / s))."""
    amax = x.float().abs().amax(dim=-1, keepdim=True)
    if scale_ub is not None:
        amax = torch.minimum(amax, scale_ub.float().reshape(()))
    fmax = torch.tensor(FP8_E4M3_MAX, dtype=torch.float32)
    min_s = torch.tensor(1.0, dtype=torch.float32) / (fmax * torch.tensor(512.0, dtype=torch.float32))
    s = torch.maximum(amax / fmax, min_s)
    return clamp_to_e4m3(x.float() / s), s


def scaled_mm(a_fp8, b_fp8, scale_a, scale_b, out_dtype, bias=None):
    """tests/kernels/utils.py:1130-1170 baseline_scaled_mm: fp32 mm((sa*a),(sb*b)) -> out dtype (+bias).
    a [M,K], b [K,N]; scales broadcast ([M,1]|[1], [1,N]|[1])."""
    sa = scale_a.float().reshape(-1, 1) if scale_a.numel() > 1 else scale_a.float().reshape(1, 1)
    sb = scale_b.float().reshape(1, -1) if scale_b.numel() > 1 else scale_b.float().reshape(1, 1)
    out = torch.mm(sa * a_fp8.float(), sb * b_fp8.float()).to(out_dtype)
    if bias is not None:
        out = out + bias
    return out


# ------------------------------------------------------------------------------ NVFP4
E2M1_VALUES = torch.tensor([0.0, 0.5, 1.0, 1.5, 2.0, 3.0, 4.0, 6.0], dtype=torch.float32)


def cast_to_e2m1_codes(x: torch.Tensor) -> torch.Tensor:
    """Round-to-nearest-even onto the e2m1 grid, saturating at 6 (cvt.rn.satfinite.e2m1x2.f32;
    thresholds as tests/kernels/quantization/test_nvfp4_quant.py:84-95). Returns 4-bit codes."""
    sign = (x < 0) | ((x == 0) & (torch.signbit(x)))
    a = x.abs()
    code = torch.zeros_like(a, dtype=torch.uint8)
    code[(a > 0.25) & (a < 0.75)] = 1
    code[(a >= 0.75) & (a <= 1.25)] = 2
    code[(a > 1.25) & (a < 1.75)] = 3
    code[(a >= 1.75) & (a <= 2.5)] = 4
    code[(a > 2.5) & (a < 3.5)] = 5
    code[(a >= 3.5) & (a <= 5.0)] = 6
    code[a > 5.0] = 7
    return code | (sign.to(torch.uint8) << 3)


def swizzle_sf(sf: torch.Tensor) -> torch.Tensor:
    """[M, K/16] uint8 -> padded [ru(M,128), ru(K/16,4)] in 128x4 tensor-core order
    (vllm/model_executor/layers/quantization/utils/nvfp4_utils.py:38-53; device-side
    offset nvfp4_utils.cuh:182-197)."""
    M, K = sf.shape
    Mp, Kp = (M + 127) // 128 * 128, (K + 3) // 4 * 4
    p = torch.zeros((Mp, Kp), dtype=sf.dtype)
    p[:M, :K] = sf
    return p.reshape(Mp // 128, 4, 32, Kp // 4, 4).permute(0, 3, 2, 1, 4).contiguous().reshape(Mp, Kp)


def unswizzle_sf(sw: torch.Tensor, m: int, k_sf: int) -> torch.Tensor:
    """Inverse (tests/kernels/quantization/nvfp4_utils.py:16-23, test_nvfp4_quant.py:122-131)."""
    Mp, Kp = (m + 127) // 128 * 128, (k_sf + 3) // 4 * 4
    t = sw.reshape(Mp // 128, Kp // 4, 32, 4, 4).permute(0, 3, 2, 1, 4).contiguous().reshape(Mp, Kp)
    return t[:m, :k_sf]


def scaled_fp4_quant(x: torch.Tensor, global_scale: torch.Tensor, swizzled: bool = True, padded_n=None):
    """tests/kernels/quantization/test_nvfp4_quant.py:107-119 (ref_nvfp4_quant) restated to
    produce the kernel's storage format: packed codes uint8 [m, n/2] (even element in the
    low nibble, :74-81) and e4m3 SF bytes (swizzled or row-major)."""
    m, n = x.shape
    out_n = n if padded_n is None else padded_n
    xf = torch.zeros((m, out_n), dtype=torch.float32)
    xf[:, :n] = x.float()
    blk = xf.reshape(m, out_n // 16, 16)
    gs = global_scale.float().reshape(())
    vec_max = blk.abs().amax(dim=-1, keepdim=True)
    sf = gs * (vec_max * (1.0 / 6.0))
    sf8 = to_fp8(sf)
    sf = sf8.float()
    rgs = torch.where(gs == 0, torch.zeros(()), 1.0 / gs)
    denom = sf * rgs
    oscale = torch.where(denom == 0, torch.zeros_like(denom), 1.0 / denom)
    codes = cast_to_e2m1_codes(torch.clamp(blk * oscale, -6.0, 6.0)).reshape(m, out_n)
    packed = codes[:, 0::2] | (codes[:, 1::2] << 4)
    sfb = sf8.view(torch.uint8).reshape(m, out_n // 16)
    return packed, (swizzle_sf(sfb) if swizzled else sfb)


def dequant_nvfp4(packed: torch.Tensor, sf_linear_u8: torch.Tensor, global_scale) -> torch.Tensor:
    """tests/kernels/quantization/nvfp4_utils.py:38-67: codes * sf / global_scale."""
    m, half = packed.shape
    lo, hi = packed & 0xF, packed >> 4
    codes = torch.stack([lo, hi], dim=-1).reshape(m, half * 2).long()
    vals = E2M1_VALUES[codes & 7] * torch.where((codes & 8) != 0, -1.0, 1.0)
    sf = sf_linear_u8.view(torch.float8_e4m3fn).float()
    return (vals.reshape(m, -1, 16) * sf[:, :, None]).reshape(m, -1) / float(global_scale)


def scaled_fp4_mm(a_packed, b_packed, sfa_swizzled, sfb_swizzled, alpha, out_dtype):
    """tests/kernels/quantization/test_nvfp4_scaled_mm.py:27-50: dequantise both operands
    (global scales folded into alpha = 1/(gs_a*gs_b)) and matmul in fp32."""
    m, k = a_packed.shape[0], a_packed.shape[1] * 2
    n = b_packed.shape[0]
    sa = unswizzle_sf(sfa_swizzled.view(torch.uint8).reshape(-1, (k // 16 + 3) // 4 * 4), m, k // 16)
    sb = unswizzle_sf(sfb_swizzled.view(torch.uint8).reshape(-1, (k // 16 + 3) // 4 * 4), n, k // 16)
    a = dequant_nvfp4(a_packed, sa, 1.0)
    b = dequant_nvfp4(b_packed, sb, 1.0)
    return (torch.mm(a, b.t()) * float(alpha)).to(out_dtype)


# ------------------------------------------------------------- element-wise companions
def rms_norm(x, w, eps, residual=None):
    """csrc/libtorch_stable/layernorm_kernels.cu: fp32 math; fused add rounds the sum to the
    activation dtype first and returns it as the new residual."""
    if residual is not None:
        x = (x.float() + residual.float()).to(x.dtype)
        residual = x
    xf = x.float()
    inv = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    out = ((xf * inv).to(x.dtype).float() * w.float()).to(x.dtype)
    return (out, residual) if residual is not None else out


def silu_and_mul(x):
    """csrc/libtorch_stable/activation_kernels.cu: silu(x[..., :d]) (rounded) * x[..., d:]."""
    d = x.shape[-1] // 2
    g = x[..., :d].float()
    s = (g / (1.0 + torch.exp(-g))).to(x.dtype)
    return (s.float() * x[..., d:].float()).to(x.dtype)


def rotary_embedding_neox(positions, q, k, head_size, cos_sin_cache):
    """csrc/libtorch_stable/pos_encoding_kernels.cu (neox): rotate halves with cos|sin rows."""
    rot = cos_sin_cache.shape[1]
    half = rot // 2
    cs = cos_sin_cache[positions.long()].float()
    c, s = cs[:, None, :half], cs[:, None, half:]

    def rot_one(t):
        T = t.shape[0]
        v = t.reshape(T, -1, head_size).float().clone()
        x, y = v[..., :half].clone(), v[..., half:rot].clone()
        v[..., :half] = x * c - y * s
        v[..., half:rot] = y * c + x * s
        return v.to(t.dtype).reshape(t.shape)

    return rot_one(q), rot_one(k)


def rms_norm_dynamic_per_token_quant(x, w, eps, scale_ub=None, residual=None):
    """csrc/libtorch_stable/quantization/fused_kernels/fused_layernorm_dynamic_per_token_quant.cu:50-91 with
    layernorm_utils.cuh (compute_rms, compute_dynamic_per_token_scales, norm_and_quant): the norm sees the
    UNROUNDED fp32 sum input + residual (residual <- T(sum)); y = T(T(x * rms) * w); s = max(min(amax|y|, ub) / 448,
    1/(448*512)); q = e4m3(y / s).  Returns (q, s [M,1], new_residual | None)."""
    xf = x.float()
    new_res = None
    if residual is not None:
        xf = xf + residual.float()
        new_res = xf.to(x.dtype)
    rms = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    y = ((xf * rms).to(x.dtype).float() * w.float()).to(x.dtype).float()
    q, s = dynamic_per_token_scaled_fp8_quant(y, scale_ub)
    return q, s, new_res


# ------------------------------------------------------------- fused NVFP4 producers
def silu_and_mul_nvfp4_quant(x, global_scale):
    """csrc/libtorch_stable/quantization/fp4/activation_nvfp4_quant_fusion_kernels.cu:39-113 with
    nvfp4_utils.cuh:304-329 (compute_silu_mul): silu(gate) * up is formed in fp32 and rounded to the activation
    dtype ONCE (one rounding fewer than silu_and_mul followed by scaled_fp4_quant), then quantised like
    scaled_fp4_quant with the swizzled scale layout.  Returns (packed codes, swizzled scale bytes)."""
    d = x.shape[-1] // 2
    g, u = x[..., :d].float(), x[..., d:].float()
    act = ((g / (1.0 + torch.exp(-g))) * u).to(x.dtype)
    return scaled_fp4_quant(act, global_scale, swizzled=True)


def rms_norm_nvfp4_quant(x, w, eps, global_scale, residual=None):
    """No reference kernel: the reference's NVFP4 path runs fused_add_rms_norm (layernorm_kernels.cu) and then
    scaled_fp4_quant (nvfp4_quant_kernels.cu); this is that composition.  Returns (packed, swizzled scales,
    new_residual | None)."""
    if residual is not None:
        y, new_res = rms_norm(x, w, eps, residual=residual)
    else:
        y, new_res = rms_norm(x, w, eps), None
    packed, sf = scaled_fp4_quant(y, global_scale, swizzled=True)
    return packed, sf, new_res
