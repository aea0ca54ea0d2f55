"""GPU: the vLLM-facing classes (attention backend impl, linear kernel classes, torch.ops.b200.*)
against the oracle AND against the INSTALLED vLLM wheel's own CUDA kernels on identical inputs
(on-GPU oracle = installed `_C` ops)."""
import types

import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu
vllm = pytest.importorskip("vllm")
DEV = "cuda:0"
ATOL, RTOL = 2e-2, 1e-2


@pytest.fixture(scope="module")
def plugin(cuda_lib):
    from vllm_b200 import plugin

    plugin.register()
    return plugin


def _layer(k=1.0, v=1.0):
    return types.SimpleNamespace(_k_scale=torch.tensor(k, device=DEV), _v_scale=torch.tensor(v, device=DEV),
                                 _q_scale=torch.tensor(1.0, device=DEV))


def _cfg(hq, hkv, d, max_len=2048, max_tokens=256):
    mc = types.SimpleNamespace(max_model_len=max_len, get_num_attention_heads=lambda pc: hq,
                               get_num_kv_heads=lambda pc: hkv, get_head_size=lambda: d)
    sc = types.SimpleNamespace(max_num_batched_tokens=max_tokens, max_num_seqs=64)
    return types.SimpleNamespace(model_config=mc, parallel_config=None, scheduler_config=sc)


@pytest.mark.parametrize("kv_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("mixed", [False, True])
def test_backend_impl_against_oracle(plugin, kv_dtype, mixed):
    """A scheduler step as vLLM issues it: do_kv_cache_update for the new tokens, then forward."""
    from vllm.v1.attention.backend import CommonAttentionMetadata
    from vllm_b200.attention_backend import B200AttentionBackend, B200AttentionImpl, B200AttentionMetadataBuilder

    Hq, Hkv, D, bs, nb = 16, 4, 128, 16, 200
    g = torch.Generator().manual_seed(5)
    if mixed:  # decode, prefill from scratch, chunked prefill with context, decode
        q_lens, ctx = [1, 37, 20, 1], [300, 0, 75, 16]
    else:
        q_lens, ctx = [1, 1, 1, 1, 1], [511, 1, 0, 130, 47]
    R, T = len(q_lens), sum(q_lens)
    seq_lens = [c + q for c, q in zip(ctx, q_lens)]
    max_blocks = (max(seq_lens) + bs - 1) // bs + 1
    bt = (torch.randperm(nb - 1, generator=g)[: R * max_blocks] + 1).reshape(R, max_blocks).to(torch.int32)
    shape = B200AttentionBackend.get_kv_cache_shape(nb, bs, Hkv, D, kv_dtype)
    ks, vs = (0.07, 0.11) if kv_dtype == "fp8" else (1.0, 1.0)
    kv_f = (torch.rand(shape, generator=g) * 2 - 1)
    if kv_dtype == "fp8":
        kv_ref = O.to_fp8(kv_f / ks).view(torch.uint8)
    else:
        kv_ref = (kv_f * D ** -0.5).to(torch.bfloat16)
    kv_dev = kv_ref.to(DEV)
    q = torch.randn(T, Hq, D, generator=g).to(torch.bfloat16)
    k_new = (torch.randn(T, Hkv, D, generator=g) * 0.5).to(torch.bfloat16)
    v_new = (torch.randn(T, Hkv, D, generator=g) * 0.5).to(torch.bfloat16)
    qsl = torch.tensor([0] + list(torch.tensor(q_lens).cumsum(0)), dtype=torch.int32)
    pos = torch.cat([torch.arange(c, c + n) for c, n in zip(ctx, q_lens)])
    req = torch.repeat_interleave(torch.arange(R), torch.tensor(q_lens))
    slots = bt[req, pos // bs].long() * bs + pos % bs

    layer = _layer(ks, vs)
    impl = B200AttentionImpl(Hq, D, D ** -0.5, Hkv, None, None, kv_dtype)
    builder = B200AttentionMetadataBuilder(types.SimpleNamespace(block_size=bs), ["l"], _cfg(Hq, Hkv, D), torch.device(DEV))
    cm = CommonAttentionMetadata(
        query_start_loc=qsl.to(DEV), query_start_loc_cpu=qsl, seq_lens=torch.tensor(seq_lens, dtype=torch.int32, device=DEV),
        num_reqs=R, num_actual_tokens=T, max_query_len=max(q_lens), max_seq_len=max(seq_lens),
        block_table_tensor=bt.to(DEV), slot_mapping=slots.to(DEV))
    md = builder.build(0, cm)
    # padded inputs like the runner's persistent buffers: rows past num_actual_tokens must stay untouched
    pad = 3
    out = torch.full((T + pad, Hq * D), 7.0, dtype=torch.bfloat16, device=DEV)
    qd = torch.cat([q, torch.zeros(pad, Hq, D, dtype=q.dtype)]).to(DEV).view(T + pad, Hq * D)
    impl.do_kv_cache_update(layer, k_new.to(DEV), v_new.to(DEV), kv_dev, md.slot_mapping)
    ret = impl.forward(layer, qd, None, None, kv_dev, md, output=out)
    assert ret is out and torch.all(out[T:] == 7.0)

    # oracle: same cache write (bit-exact), then one causal single-query attention per token
    kc_ref, vc_ref = kv_ref[0].clone(), kv_ref[1].clone()
    if kv_dtype == "fp8":
        kc_ref, vc_ref = kc_ref.view(torch.float8_e4m3fn), vc_ref.view(torch.float8_e4m3fn)
    O.reshape_and_cache_flash(k_new, v_new, kc_ref, vc_ref, slots, kv_dtype, torch.tensor([ks]), torch.tensor([vs]))
    assert torch.equal(kv_dev[0].cpu().view(torch.uint8), kc_ref.view(torch.uint8))
    assert torch.equal(kv_dev[1].cpu().view(torch.uint8), vc_ref.view(torch.uint8))
    ref = O.paged_attention_decode(q, kc_ref, vc_ref, bt[req], (pos + 1).to(torch.int32), D ** -0.5, kv_dtype, ks, vs)
    torch.testing.assert_close(out[:T].cpu().float().view(T, Hq, D), ref, atol=ATOL, rtol=RTOL)


def test_backend_impl_against_installed_flash_attn(plugin):
    """Same decode step through vLLM's own FlashAttention op (what FlashAttentionImpl.forward calls,
    flash_attn.py:1085-1110) on the same paged cache."""
    try:
        from vllm.v1.attention.backends.fa_utils import flash_attn_varlen_func
    except Exception as e:  # pragma: no cover
        pytest.skip(f"installed vllm flash-attn op unavailable: {e}")
    from vllm_b200.attention_backend import B200AttentionImpl, B200AttentionMetadata

    Hq, Hkv, D, bs, nb, S = 32, 8, 128, 16, 600, 6
    g = torch.Generator(device=DEV).manual_seed(3)
    lens = torch.tensor([1000, 17, 512, 333, 1, 64], dtype=torch.int32, device=DEV)
    mb = 64
    bt = (torch.randperm(nb - 1, generator=g, device=DEV)[: S * mb] + 1).reshape(S, mb).to(torch.int32)
    kv = ((torch.rand(2, nb, bs, Hkv, D, device=DEV, generator=g) * 2 - 1) * D ** -0.5).to(torch.bfloat16)
    q = torch.randn(S, Hq, D, device=DEV, generator=g).to(torch.bfloat16)
    try:
        ref = flash_attn_varlen_func(q=q, k=kv[0], v=kv[1], cu_seqlens_q=torch.arange(S + 1, dtype=torch.int32, device=DEV),
                                     max_seqlen_q=1, seqused_k=lens, max_seqlen_k=1000, softmax_scale=D ** -0.5,
                                     causal=True, block_table=bt)
    except Exception as e:  # pragma: no cover  (FA build without sm_100 support)
        pytest.skip(f"installed flash_attn_varlen_func does not run here: {e}")
    impl = B200AttentionImpl(Hq, D, D ** -0.5, Hkv, None, None, "auto")
    md = B200AttentionMetadata(num_actual_tokens=S, max_query_len=1, max_seq_len=1000,
                               query_start_loc=torch.arange(S + 1, dtype=torch.int32, device=DEV), seq_lens=lens,
                               block_table=bt, slot_mapping=torch.zeros(S, dtype=torch.int64, device=DEV), num_splits=0)
    out = torch.empty(S, Hq, D, dtype=torch.bfloat16, device=DEV)
    impl.forward(_layer(), q, None, None, kv, md, output=out)
    torch.testing.assert_close(out.float(), ref.float(), atol=ATOL, rtol=RTOL)


def _fp8_kernel(act_key, w_key, n, k):
    from vllm.model_executor.kernels.linear import FP8ScaledMMLinearLayerConfig
    from vllm_b200.linear_kernels import B200FP8ScaledMMLinearKernel

    from vllm.config import VllmConfig, set_current_vllm_config

    c = FP8ScaledMMLinearLayerConfig(weight_quant_key=w_key, activation_quant_key=act_key, weight_shape=(n, k),
                                     input_dtype=torch.bfloat16, out_dtype=torch.bfloat16)
    assert B200FP8ScaledMMLinearKernel.is_supported()[0] and B200FP8ScaledMMLinearKernel.can_implement(c)[0]
    with set_current_vllm_config(VllmConfig()):  # the base class builds a QuantFP8 CustomOp
        return B200FP8ScaledMMLinearKernel(c, ["weight", "weight_scale", "input_scale", "input_scale_ub"])


@pytest.mark.parametrize("own_quant", [False, True])
@pytest.mark.parametrize("static_act", [False, True])
@pytest.mark.parametrize("m", [1, 128, 300])
def test_fp8_linear_kernel_against_installed_cutlass(plugin, static_act, m, own_quant, monkeypatch):
    """B200FP8ScaledMMLinearKernel.apply_weights == vLLM's quant + cutlass_scaled_mm on the same layer."""
    from vllm import _custom_ops as vops
    from vllm.model_executor.layers.quantization.utils import quant_utils as Q

    n, k = 768, 1024
    g = torch.Generator(device=DEV).manual_seed(m)
    w = (torch.randn(n, k, device=DEV, generator=g) * 0.05).to(torch.bfloat16)
    x = torch.randn(m, k, device=DEV, generator=g).to(torch.bfloat16)
    bias = torch.randn(n, device=DEV, generator=g).to(torch.bfloat16)
    wq, ws = vops.scaled_fp8_quant(w)  # per-tensor, as Fp8PerTensorOnlineLinearMethod does (online/fp8.py:207-216)
    layer = types.SimpleNamespace(weight=wq.t(), weight_scale=ws,
                                  input_scale=torch.tensor([0.02], device=DEV) if static_act else None, input_scale_ub=None)
    kern = _fp8_kernel(Q.kFp8StaticTensorSym if static_act else Q.kFp8DynamicTokenSym, Q.kFp8StaticTensorSym, n, k)
    kern.process_weights_after_loading(layer)
    # own_quant False: activation quant by vLLM's QuantFP8 (the default, fusible by Inductor); True: our quant ops
    from vllm_b200 import linear_kernels

    monkeypatch.setattr(linear_kernels, "_OWN_QUANT", own_quant)
    y = kern.apply_weights(layer, x, bias)
    if not own_quant:
        # the reference quantises with the SAME QuantFP8 object the kernel class inherited (in eager mode its native
        # torch formula, which rounds a few elements differently from the _C kernel): only the GEMM differs
        xq, xs = kern.quant_fp8(x, layer.input_scale, None)
    elif static_act:
        xq, xs = vops.scaled_fp8_quant(x, layer.input_scale)
    else:
        xq, xs = vops.scaled_fp8_quant(x, use_per_token_if_dynamic=True)
    try:
        ref = vops.cutlass_scaled_mm(xq, wq.t(), scale_a=xs, scale_b=ws, out_dtype=torch.bfloat16, bias=bias)
    except Exception as e:  # pragma: no cover
        pytest.skip(f"installed cutlass_scaled_mm does not run here: {e}")
    assert y.shape == (m, n)
    torch.testing.assert_close(y.float(), ref.float(), atol=2e-2, rtol=2e-2)


def test_nvfp4_linear_kernel_against_installed_cutlass(plugin):
    from vllm import _custom_ops as vops
    from vllm.model_executor.kernels.linear import NvFp4LinearLayerConfig
    from vllm_b200 import ops
    from vllm_b200.linear_kernels import B200NvFp4LinearKernel

    m, n, k = 130, 512, 1024
    g = torch.Generator(device=DEV).manual_seed(0)
    w = (torch.randn(n, k, device=DEV, generator=g) * 0.1).to(torch.bfloat16)
    x = torch.randn(m, k, device=DEV, generator=g).to(torch.bfloat16)
    w_gs = (448.0 * 6.0 / w.abs().max().float()).reshape(1)
    x_gs = (448.0 * 6.0 / x.abs().max().float()).reshape(1)
    wq, wsf = ops.scaled_fp4_quant(w, w_gs, is_sf_swizzled_layout=False)  # ModelOpt checkpoint layout: linear SFs
    layer = types.SimpleNamespace(weight=wq, weight_scale=wsf.view(torch.float8_e4m3fn)[:n, : k // 16].contiguous(),
                                  alpha=(1.0 / (w_gs * x_gs)).float(), input_global_scale_inv=x_gs,
                                  output_size_per_partition=n)
    kern = B200NvFp4LinearKernel(NvFp4LinearLayerConfig())
    kern.process_weights_after_loading(layer)
    y = kern.apply_weights(layer, x, None)
    try:
        xq, xsf = vops.scaled_fp4_quant(x, x_gs)
        ref = vops.cutlass_scaled_fp4_mm(xq, layer.weight, xsf, layer.weight_scale, layer.alpha, torch.bfloat16)
    except Exception as e:  # pragma: no cover
        pytest.skip(f"installed nvfp4 kernels do not run here: {e}")
    torch.testing.assert_close(y.float(), ref.float(), atol=1e-1, rtol=1e-1)  # test_nvfp4_scaled_mm.py:100
    dense = x.float() @ w.float().t()
    assert (y.float() - dense).abs().mean() < 0.15 * dense.abs().mean()


def test_torch_ops_b200_run_on_cuda(plugin):
    """The torch.ops surface (B3) end to end: cache write + fp8 quant + scaled_mm through torch.ops.b200."""
    T, H, D, bs, nb = 9, 4, 128, 16, 8
    g = torch.Generator().manual_seed(0)
    key, val = torch.randn(T, H, D, generator=g).to(torch.bfloat16), torch.randn(T, H, D, generator=g).to(torch.bfloat16)
    kc, vc = torch.zeros(nb, bs, H, D, dtype=torch.bfloat16), torch.zeros(nb, bs, H, D, dtype=torch.bfloat16)
    slots = torch.randperm(nb * bs, generator=g)[:T]
    kcd, vcd = kc.to(DEV), vc.to(DEV)
    one = torch.ones(1, device=DEV)
    torch.ops.b200.reshape_and_cache_flash(key.to(DEV), val.to(DEV), kcd, vcd, slots.to(DEV), "auto", one, one)
    O.reshape_and_cache_flash(key, val, kc, vc, slots, "auto", torch.ones(1), torch.ones(1))
    assert torch.equal(kcd.cpu(), kc) and torch.equal(vcd.cpu(), vc)

    x = torch.randn(33, 256, generator=g).to(torch.bfloat16)
    res = torch.empty(33, 256, dtype=torch.float8_e4m3fn, device=DEV)
    sc = torch.empty(33, 1, device=DEV)
    torch.ops.b200.dynamic_per_token_scaled_fp8_quant(res, x.to(DEV), sc, None)
    rq, rs = O.dynamic_per_token_scaled_fp8_quant(x)
    assert torch.equal(sc.cpu(), rs) and torch.equal(res.cpu().view(torch.uint8), rq.view(torch.uint8))
    w = (torch.randn(64, 256, generator=g) * 0.1).to(torch.bfloat16)
    wq = torch.empty(64, 256, dtype=torch.float8_e4m3fn, device=DEV)
    wsc = torch.empty(1, device=DEV)
    torch.ops.b200.dynamic_scaled_fp8_quant(wq, w.to(DEV), wsc)
    out = torch.empty(33, 64, dtype=torch.bfloat16, device=DEV)
    torch.ops.b200.cutlass_scaled_mm(out, res, wq.t(), sc, wsc, None)
    ref = O.scaled_mm(rq, wq.cpu().t(), rs, wsc.cpu(), torch.float32)
    torch.testing.assert_close(out.cpu().float(), ref, atol=2e-2, rtol=1e-2)


def test_torch_ops_b200_fused_elementwise(plugin):
    """The norm / activation / rope ops and their fused-quant forms through torch.ops.b200 (the reference's schemas,
    torch_bindings.cpp:251-253,370-377,394-396,418-421,582) == the same kernels through vllm_b200.ops."""
    from vllm_b200 import ops

    g = torch.Generator().manual_seed(1)
    M, Hd = 19, 512
    x = torch.randn(M, Hd, generator=g).to(torch.bfloat16).to(DEV)
    w = (torch.rand(Hd, generator=g) + 0.5).to(torch.bfloat16).to(DEV)
    r = torch.randn(M, Hd, generator=g).to(torch.bfloat16).to(DEV)
    out, ref = torch.empty_like(x), torch.empty_like(x)
    torch.ops.b200.rms_norm(out, x, w, 1e-5)
    ops.rms_norm(ref, x, w, 1e-5)
    assert torch.equal(out, ref)
    x1, r1, x2, r2 = x.clone(), r.clone(), x.clone(), r.clone()
    torch.ops.b200.fused_add_rms_norm(x1, r1, w, 1e-5)
    ops.fused_add_rms_norm(x2, r2, w, 1e-5)
    assert torch.equal(x1, x2) and torch.equal(r1, r2)
    q = torch.empty(M, Hd, dtype=torch.float8_e4m3fn, device=DEV)
    sc = torch.empty(M, 1, device=DEV)
    r3, r4 = r.clone(), r.clone()
    torch.ops.b200.rms_norm_dynamic_per_token_quant(q, x, w, sc, 1e-5, None, r3)
    q2, sc2 = ops.rms_norm_dynamic_per_token_quant(x, w, 1e-5, residual=r4)
    assert torch.equal(q.view(torch.uint8), q2.view(torch.uint8)) and torch.equal(sc, sc2) and torch.equal(r3, r4)
    gu = torch.randn(M, 2 * Hd, generator=g).to(torch.bfloat16).to(DEV)
    a1, a2 = torch.empty(M, Hd, dtype=torch.bfloat16, device=DEV), torch.empty(M, Hd, dtype=torch.bfloat16, device=DEV)
    torch.ops.b200.silu_and_mul(a1, gu)
    ops.silu_and_mul(a2, gu)
    assert torch.equal(a1, a2)
    gs = torch.tensor([448.0 * 6.0 / 10.0], device=DEV)
    c1, c2 = (torch.empty(M, Hd // 2, dtype=torch.uint8, device=DEV) for _ in range(2))
    s1, s2 = (torch.zeros(128, Hd // 64, dtype=torch.int32, device=DEV) for _ in range(2))
    torch.ops.b200.silu_and_mul_nvfp4_quant(c1, s1, gu, gs)
    ops.silu_and_mul_nvfp4_quant(c2, s2, gu, gs)
    assert torch.equal(c1, c2) and torch.equal(s1, s2)
    T, Hq, D = 7, 4, 128
    pos = torch.arange(T, device=DEV)
    cs = torch.randn(64, D, generator=g).to(torch.bfloat16).to(DEV)
    qk = torch.randn(T, 2 * Hq * D, generator=g).to(torch.bfloat16).to(DEV)
    qa, ka = qk.clone()[:, :Hq * D], qk.clone()[:, Hq * D:]
    qb, kb = qk.clone()[:, :Hq * D], qk.clone()[:, Hq * D:]
    torch.ops.b200.rotary_embedding(pos, qa, ka, D, cs, True)
    ops.rotary_embedding(pos, qb, kb, D, cs, True)
    assert torch.equal(qa, qb) and torch.equal(ka, kb) and not torch.equal(qa, qk[:, :Hq * D])
    with pytest.raises(RuntimeError):
        torch.ops.b200.rotary_embedding(pos, qa, ka, D, cs, False)


@pytest.mark.parametrize("kv_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("is_neox", [True, False])
def test_fused_rope_kvcache_hook_matches_unfused(plugin, kv_dtype, is_neox):
    """AttentionImpl.do_rope_and_kv_cache_update (vllm/v1/attention/backend.py:932-980) == the reference's own
    rotary_embedding op followed by do_kv_cache_update: q, k and both cache planes bit-identical, pad slots untouched,
    q/k/v passed as slices of the fused qkv buffer the way the model does."""
    from vllm import _custom_ops as vops
    from vllm_b200.attention_backend import B200AttentionBackend, B200AttentionImpl

    Hq, Hkv, D, bs, nb, T = 16, 4, 128, 16, 40, 23
    g = torch.Generator(device=DEV).manual_seed(11)
    impl = B200AttentionImpl(Hq, D, D ** -0.5, Hkv, None, None, kv_dtype)
    assert impl.fused_rope_kvcache_supported()
    layer = _layer(0.05, 0.07) if kv_dtype == "fp8" else _layer()
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, device=DEV, generator=g).to(torch.bfloat16)
    pos = torch.randint(0, 4000, (T,), device=DEV, generator=g)
    inv = 1.0 / (10000.0 ** (torch.arange(0, D, 2, device=DEV).float() / D))
    fr = torch.outer(torch.arange(4096, device=DEV).float(), inv)
    cos_sin = torch.cat([fr.cos(), fr.sin()], -1).to(torch.bfloat16)
    slots = torch.randperm(nb * bs, device=DEV, generator=g)[:T].to(torch.int64)
    slots[3] = -1
    shape = B200AttentionBackend.get_kv_cache_shape(nb, bs, Hkv, D, kv_dtype)
    cdt = torch.uint8 if kv_dtype == "fp8" else torch.bfloat16
    kv_a = torch.zeros(shape, dtype=cdt, device=DEV)
    kv_b = torch.zeros_like(kv_a)
    # unfused: reference rope op, then our cache write
    qkv_a = qkv.clone()
    qa, ka, va = qkv_a.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    vops.rotary_embedding(pos, qa, ka, D, cos_sin, is_neox)
    impl.do_kv_cache_update(layer, ka.view(T, Hkv, D), va.view(T, Hkv, D), kv_a, slots)
    # fused hook
    qkv_b = qkv.clone()
    qb, kb, vb = qkv_b.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    impl.do_rope_and_kv_cache_update(layer, qb, kb, vb, pos, cos_sin, is_neox, kv_b, slots)
    torch.cuda.synchronize()
    assert torch.equal(qkv_a, qkv_b), "rotated q/k differ from the reference rotary_embedding op"
    assert torch.equal(kv_a, kv_b), "cache contents differ from the unfused write"
    assert kv_b.any()
