"""The REFERENCE's GPU attention-decode kernels on the same paged cache tensors, for timing next to ours
(scripts/bench_kernels.py --what attn --ref; SURVEY 8d "reference timed beside it").  Measurement aid only:
nothing under vllm_b200/ imports this.

  flashinfer_trtllm   FlashInfer trtllm-gen decode (vllm/v1/attention/backends/flashinfer.py:19, the Blackwell default)
  flash_attn          vLLM's FlashAttention varlen op over the paged cache (flash_attn.py:1085-1110); 16-bit caches
  triton_unified      kernel_unified_attention (triton_unified_attention.py:179-681), 2D or 3D exactly as
                      TritonAttentionImpl would choose (seq_threshold_3D = 128 // num_kv_heads)
  paged_attention_v2  the legacy CUDA kernel in the INSTALLED wheel (INSTALLED:_custom_ops.py:158-205); needs the
                      legacy cache layout, built here once per cache (not timed)

Every builder returns (callable, note) or raises; the callable writes into its own output tensor.
"""
from __future__ import annotations

import torch


def build_flashinfer_trtllm(q, kc, vc, bt, sl, scale, max_len, kv, ks, vs):
    from flashinfer.decode import trtllm_batch_decode_with_kv_cache

    dev = q.device
    ws = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
    out = torch.empty_like(q)
    note = "q bf16"
    if kv != "auto":
        k, v = kc.view(torch.float8_e4m3fn), vc.view(torch.float8_e4m3fn)
        # vLLM quantises the query to fp8 for trtllm-gen with an fp8 cache (flashinfer.py, q_data_type = kv dtype)
        qq = q.to(torch.float8_e4m3fn)
        note = "q quantised to fp8 (as vLLM's FlashInfer backend does for an fp8 cache); the cast is not timed"
        b1, b2 = scale * float(ks), float(vs)
    else:
        k, v, qq, b1, b2 = kc, vc, q, scale, 1.0

    def run():
        trtllm_batch_decode_with_kv_cache(query=qq, kv_cache=(k, v), workspace_buffer=ws, block_tables=bt, seq_lens=sl,
                                          max_seq_len=max_len, bmm1_scale=b1, bmm2_scale=b2, out=out, kv_layout="NHD")

    run()
    torch.cuda.synchronize()
    return run, out, note


def build_flash_attn(q, kc, vc, bt, sl, scale, max_len, kv, ks, vs):
    if kv != "auto":
        raise RuntimeError("FA2 path has no fp8-cache support on sm_100")
    from vllm.v1.attention.backends.fa_utils import flash_attn_varlen_func

    S = q.shape[0]
    cu = torch.arange(S + 1, dtype=torch.int32, device=q.device)
    out = torch.empty_like(q)

    def run():
        flash_attn_varlen_func(q=q, k=kc, v=vc, out=out, cu_seqlens_q=cu, max_seqlen_q=1, seqused_k=sl,
                               max_seqlen_k=max_len, softmax_scale=scale, causal=True, block_table=bt)

    run()
    torch.cuda.synchronize()
    return run, out, "vllm_flash_attn varlen over block_table"


def build_triton_unified(q, kc, vc, bt, sl, scale, max_len, kv, ks, vs):
    from vllm.v1.attention.ops.triton_unified_attention import unified_attention

    S, Hq, D = q.shape
    Hkv = kc.shape[2]
    dev = q.device
    cu = torch.arange(S + 1, dtype=torch.int32, device=dev)
    out = torch.empty_like(q)
    thr = max(1, 128 // Hkv)
    dpad = 1 << (D - 1).bit_length()
    seg_o = torch.empty(thr, Hq, 16, dpad, dtype=torch.float32, device=dev)
    seg_m = torch.empty(thr, Hq, 16, dtype=torch.float32, device=dev)
    seg_e = torch.empty(thr, Hq, 16, dtype=torch.float32, device=dev)
    kw = {}
    if kv != "auto":
        k, v = kc.view(torch.float8_e4m3fn), vc.view(torch.float8_e4m3fn)
        kd, vd = ks.expand(S, Hkv), vs.expand(S, Hkv)
        try:
            from vllm.v1.kv_cache_interface import KVQuantMode

            kw["kv_quant_mode"] = KVQuantMode.FP8_PER_TENSOR
        except Exception:
            pass
    else:
        k, v, kd, vd = kc, vc, None, None

    def run():
        unified_attention(q=q, k=k, v=v, out=out, cu_seqlens_q=cu, max_seqlen_q=1, seqused_k=sl, max_seqlen_k=max_len,
                          softmax_scale=scale, causal=True, window_size=(-1, -1), block_table=bt, softcap=0,
                          q_descale=None, k_descale=kd, v_descale=vd, seq_threshold_3D=thr, num_par_softmax_segments=16,
                          softmax_segm_output=seg_o, softmax_segm_max=seg_m, softmax_segm_expsum=seg_e, **kw)

    run()
    torch.cuda.synchronize()
    return run, out, f"{'3D (split-KV, 16 segments)' if S <= thr else '2D (no split)'} as TritonAttentionImpl selects"


def build_paged_attention_v2(q, kc, vc, bt, sl, scale, max_len, kv, ks, vs):
    import vllm._C  # noqa: F401

    S, Hq, D = q.shape
    nb, bs, Hkv, _ = kc.shape
    x = 16 // kc.element_size()
    k_leg = kc.permute(0, 2, 3, 1).reshape(nb, Hkv, D // x, x, bs).permute(0, 1, 2, 4, 3).contiguous()
    v_leg = vc.permute(0, 2, 3, 1).contiguous()
    P = (max_len + 511) // 512
    out = torch.empty_like(q)
    es = torch.empty(S, Hq, P, dtype=torch.float32, device=q.device)
    ml = torch.empty_like(es)
    tmp = torch.empty(S, Hq, P, D, dtype=q.dtype, device=q.device)
    one = torch.ones(1, device=q.device)
    kss, vss = (ks, vs) if kv != "auto" else (one, one)

    def run():
        torch.ops._C.paged_attention_v2(out, es, ml, tmp, q, k_leg, v_leg, Hkv, scale, bt, sl, bs, max_len, None, kv, kss,
                                        vss, 0, 0, 0, 64, 0)

    run()
    torch.cuda.synchronize()
    return run, out, "legacy layout copy of the cache (built untimed); one CTA per q-head: KV re-read G times"


BUILDERS = {
    "flashinfer_trtllm": build_flashinfer_trtllm,
    "flash_attn": build_flash_attn,
    "triton_unified": build_triton_unified,
    "paged_attention_v2": build_paged_attention_v2,
}
