"""b200_paged_attention_decode_static_scratch: the allocation-free variant of the paged-attention decode entry
point (split-KV partials in a file-scope `static __device__` array, include/b200_vllm.h).  No reference
counterpart.

Called through the C ABI (ctypes) and compared with the existing caller-workspace entry point on the same
inputs, on two calls in a row with DIFFERENT shapes / split counts (the second call finds the first call's
partials in the shared array and must not be affected by them).  Collected last on purpose (file name), passes
alone in a fresh process and after the other tests.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

ATOL, RTOL = 2e-2, 1e-2  # the attention tolerance used throughout tests/
DEV = "cuda:0"


def _params(_lib, out, q, kc, vc, bt, sl, scale, max_len, kv, ks, vs, splits, workspace):
    return _lib.PagedDecodeParams(
        out=out.data_ptr(), query=q.data_ptr(), key_cache=kc.data_ptr(), value_cache=vc.data_ptr(),
        block_table=bt.data_ptr(), seq_lens=sl.data_ptr(), k_scale=None if ks is None else ks.data_ptr(),
        v_scale=None if vs is None else vs.data_ptr(), out_lse=None, token_to_seq=None,
        workspace=None if workspace is None else workspace.data_ptr(),
        workspace_bytes=0 if workspace is None else workspace.numel(),
        q_stride_seq=q.stride(0), q_stride_head=q.stride(1), out_stride_seq=out.stride(0), out_stride_head=out.stride(1),
        kv_block_stride=kc.stride(0), kv_page_stride=kc.stride(1), kv_head_stride=kc.stride(2),
        v_block_stride=vc.stride(0), v_page_stride=vc.stride(1), v_head_stride=vc.stride(2),
        block_table_stride=bt.stride(0), num_blocks=kc.shape[0], num_seqs=q.shape[0], num_q_heads=q.shape[1],
        num_kv_heads=kc.shape[2], head_size=q.shape[2], block_size=kc.shape[1], max_seq_len=int(max_len),
        q_dtype=_lib.DT_BF16, kv_dtype=_lib.KV_AUTO if kv == "auto" else _lib.KV_FP8_E4M3, num_splits=int(splits),
        scale=float(scale), logits_soft_cap=0.0, sliding_window=0)


def _case(seed, S, Hq, Hkv, D, bs, lens, kv):
    g = torch.Generator(device=DEV).manual_seed(seed)
    maxb = (max(lens) + bs - 1) // bs
    nb = S * maxb + 3
    if kv == "auto":
        kc = ((torch.rand(nb, bs, Hkv, D, device=DEV, generator=g) * 2 - 1) * D ** -0.5).to(torch.bfloat16)
        vc = (torch.rand(nb, bs, Hkv, D, device=DEV, generator=g) * 2 - 1).to(torch.bfloat16)
    else:
        kc = torch.randint(0, 256, (nb, bs, Hkv, D), device=DEV, generator=g, dtype=torch.uint8) & 0x77
        vc = torch.randint(0, 256, (nb, bs, Hkv, D), device=DEV, generator=g, dtype=torch.uint8) & 0x77
    bt = (torch.randperm(nb - 1, device=DEV, generator=g)[: S * maxb] + 1).reshape(S, maxb).to(torch.int32)
    q = torch.randn(S, Hq, D, device=DEV, generator=g).to(torch.bfloat16)
    sl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    return q, kc, vc, bt, sl


def test_static_scratch_matches_workspace_entry_point_twice(cuda_lib):
    from vllm_b200 import _lib

    lib = cuda_lib
    cap = int(lib.b200_paged_decode_static_scratch_bytes())
    assert cap == 32 << 20
    st = torch.cuda.current_stream(torch.device(DEV)).cuda_stream
    # two calls in a row: (bf16 cache, 6 splits, ragged with an EMPTY sequence) then (fp8 cache, 2 splits, shorter)
    calls = [
        dict(seed=1, S=12, Hq=32, Hkv=8, D=128, bs=16, lens=[2000, 17, 0, 1555, 640, 1, 2048, 900, 333, 1200, 64, 1999],
             kv="auto", splits=6),
        dict(seed=2, S=5, Hq=8, Hkv=1, D=128, bs=32, lens=[700, 31, 1024, 5, 512], kv="fp8", splits=2),
        dict(seed=3, S=12, Hq=32, Hkv=8, D=128, bs=16, lens=[100] * 12, kv="auto", splits=0),  # library-picked splits
    ]
    for i, c in enumerate(calls):
        q, kc, vc, bt, sl = _case(c["seed"], c["S"], c["Hq"], c["Hkv"], c["D"], c["bs"], c["lens"], c["kv"])
        ks = torch.tensor([0.0625], device=DEV) if c["kv"] != "auto" else None
        vs = torch.tensor([0.125], device=DEV) if c["kv"] != "auto" else None
        scale, mx = c["D"] ** -0.5, max(c["lens"])
        # existing entry point with a caller workspace (sized for the largest split count it may pick)
        ws = torch.empty(int(lib.b200_paged_decode_workspace_bytes(c["S"], c["Hq"], c["D"], 128)) + 16, dtype=torch.uint8,
                         device=DEV)
        ref = torch.full_like(q, float("nan"))
        p_ref = _params(_lib, ref, q, kc, vc, bt, sl, scale, mx, c["kv"], ks, vs, c["splits"], ws)
        _lib.check(lib.b200_paged_attention_decode(C.byref(p_ref), st), "paged_attention_decode")
        # new entry point: no workspace passed at all
        out = torch.full_like(q, float("nan"))
        p_new = _params(_lib, out, q, kc, vc, bt, sl, scale, mx, c["kv"], ks, vs, c["splits"], None)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        alloc0 = torch.cuda.memory_allocated()
        _lib.check(lib.b200_paged_attention_decode_static_scratch(C.byref(p_new), st),
                   "paged_attention_decode_static_scratch")
        torch.cuda.synchronize()
        if i > 0:  # (the very first launch may trigger lazy module loading, which is not an allocation of ours)
            assert torch.cuda.mem_get_info()[0] == free0, "the allocation-free entry point changed free device memory"
        assert torch.cuda.memory_allocated() == alloc0
        assert torch.isfinite(out.float()).all()
        torch.testing.assert_close(out.float(), ref.float(), atol=ATOL, rtol=RTOL)
        empty = sl == 0
        assert torch.equal(out[empty], torch.zeros_like(out[empty]))  # sequences of length 0 produce zeros


def test_static_scratch_rejects_oversize_before_launch(cuda_lib):
    from vllm_b200 import _lib

    lib = cuda_lib
    # 4096 seqs x 64 heads x 8 splits x (128*4+8) B = 1.09 GB of partials > 32 MiB: refused, nothing launched
    p = _lib.PagedDecodeParams(num_seqs=4096, num_q_heads=64, num_kv_heads=8, head_size=128, block_size=16,
                               max_seq_len=8192, q_dtype=_lib.DT_BF16, kv_dtype=_lib.KV_AUTO, num_splits=8)
    before = _lib.num_kernels_launched()
    rc = lib.b200_paged_attention_decode_static_scratch(C.byref(p), None)
    assert rc == _lib.ERR_WORKSPACE
    msg = lib.b200_last_error().decode()
    assert "built-in" in msg and str(32 << 20) in msg, msg
    assert _lib.num_kernels_launched() == before
    # a single split needs no scratch and is accepted up to the ordinary argument checks (null pointers here)
    p.num_splits = 1
    assert lib.b200_paged_attention_decode_static_scratch(C.byref(p), None) == _lib.ERR_INVALID_ARG
