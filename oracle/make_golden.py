"""Generate tests/golden/*.pt from the REFERENCE's own code, run in this container.

TEST INFRASTRUCTURE ONLY.  /root/reference does not exist on the GPU box, so the
vectors are committed; this script is what made them.  Sources of truth used:

  * the reference's Python test references, imported by extracting the named
    functions/constants from the files under /root/reference (module-level imports
    of those files need a GPU or the full test package, so the defs are exec'd in a
    small namespace; nothing is edited):
      tests/kernels/attention/test_triton_unified_attention.py  ref_paged_attn
      tests/kernels/attention/test_attention.py                 ref_single_query_cached_kv_attention
      tests/kernels/attention/test_merge_attn_states.py         merge_attn_states_torch
      tests/kernels/quantization/test_nvfp4_quant.py            ref_nvfp4_quant, cast_to_fp4, recover_swizzled_scales
      tests/kernels/quantization/nvfp4_utils.py                 dequantize_nvfp4_to_dtype, ...
      tests/kernels/quant_utils.py                              ref_dynamic_per_token_quant, ref_dynamic_per_tensor_fp8_quant
      tests/kernels/utils.py                                    baseline_scaled_mm
      vllm/model_executor/layers/quantization/utils/nvfp4_utils.py  swizzle_blockscale
  * the reference's CPU attention kernels compiled unmodified (oracle/_ref, see
    oracle/build_ref.py): cpu_attn_reshape_and_cache + cpu_attention_with_kv_cache.

Run:  python -m oracle.make_golden
"""
from __future__ import annotations

import ast
import math
import os
from pathlib import Path

import torch

REF = Path(os.environ.get("VLLM_REFERENCE", "/root/reference"))
OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"


def extract(path: Path, names: list[str], ns: dict) -> dict:
    """exec the top-level defs/assignments called `names` from `path` into `ns`."""
    tree = ast.parse(path.read_text())
    picked = []
    for node in tree.body:
        if isinstance(node, (ast.FunctionDef, ast.ClassDef)) and node.name in names:
            picked.append(node)
        elif isinstance(node, ast.Assign) and any(isinstance(t, ast.Name) and t.id in names for t in node.targets):
            picked.append(node)
    found = {getattr(n, "name", None) or n.targets[0].id for n in picked}
    missing = set(names) - found
    if missing:
        raise RuntimeError(f"{path}: reference no longer defines {sorted(missing)}")
    mod = ast.Module(body=picked, type_ignores=[])
    exec(compile(mod, str(path), "exec"), ns)
    return ns


def gen(seed: int) -> torch.Generator:
    g = torch.Generator()
    g.manual_seed(seed)
    return g


def make_paged_case(seed, S, Hq, Hkv, D, bs, lens, dtype, num_blocks):
    g = gen(seed)
    scale = D ** -0.5
    q = torch.randn(S, Hq, D, generator=g).to(dtype)
    # cache init as vllm/utils/torch_utils.py:621,647: uniform(-scale, scale)
    kc = (torch.rand(num_blocks, bs, Hkv, D, generator=g) * 2 - 1).mul(scale).to(dtype)
    vc = (torch.rand(num_blocks, bs, Hkv, D, generator=g) * 2 - 1).mul(scale).to(dtype)
    maxb = (max(lens) + bs - 1) // bs
    perm = torch.randperm(num_blocks - 1, generator=g)[: S * maxb] + 1  # block 0 = null block
    bt = perm.reshape(S, maxb).to(torch.int32)
    return q, kc, vc, bt, torch.tensor(lens, dtype=torch.int32), scale


def main() -> None:
    OUT.mkdir(parents=True, exist_ok=True)
    torch.set_num_threads(8)
    T = REF / "tests" / "kernels"

    # ---------------- paged attention: ref_paged_attn + legacy reference + compiled CPU kernels
    ns = {"torch": torch}
    extract(T / "attention" / "test_triton_unified_attention.py", ["ref_paged_attn"], ns)
    extract(T / "attention" / "test_attention.py", ["ref_masked_attention", "ref_single_query_cached_kv_attention"], ns)
    from . import ref_cpu

    cases = []
    specs = [
        # seed, S, Hq, Hkv, D, bs, lens, dtype
        (0, 3, 8, 2, 128, 16, [523, 37, 300], torch.bfloat16),    # (8,2) heads of test_triton_unified_attention
        (1, 4, 5, 1, 128, 16, [1, 16, 17, 300], torch.bfloat16),  # (5,1) heads of the same test
        (2, 3, 8, 2, 128, 32, [200, 33, 96], torch.bfloat16),     # Llama GQA ratio 4, block 32 (CPU kernels need %32)
        (3, 4, 4, 4, 64, 32, [130, 64, 7, 96], torch.bfloat16),    # OPT-125m style MHA, D=64 (cfg1 shape)
        (4, 2, 16, 2, 128, 16, [300, 100], torch.float16),        # G=8 like test_attention.py (64,8), fp16
    ]
    for seed, S, Hq, Hkv, D, bs, lens, dtype in specs:
        nb = S * ((max(lens) + bs - 1) // bs) + 8
        q, kc, vc, bt, sl, scale = make_paged_case(seed, S, Hq, Hkv, D, bs, lens, dtype, nb)
        out_ref = ns["ref_paged_attn"](q.clone(), kc, vc, [1] * S, lens, bt, scale)
        # legacy layout for ref_single_query_cached_kv_attention: K [B,H,D/x,bs,x], V [B,H,D,bs]
        x = 16 // q.element_size()
        k_leg = kc.permute(0, 2, 3, 1).reshape(nb, Hkv, D // x, x, bs).permute(0, 1, 2, 4, 3).contiguous()
        v_leg = vc.permute(0, 2, 3, 1).contiguous()
        out_leg = torch.empty_like(q)
        ns["ref_single_query_cached_kv_attention"](out_leg, q, Hq // Hkv, k_leg, v_leg, bt, sl, scale, None)
        case = dict(q=q, key_cache=kc, value_cache=vc, block_table=bt, seq_lens=sl, scale=scale,
                    out_ref_paged_attn=out_ref, out_legacy_ref=out_leg)
        if ref_cpu.available() and bs % 32 == 0 and dtype == torch.bfloat16:
            isa = ref_cpu.pick_isa(dtype, bs, D)
            kc_h = torch.zeros(nb, Hkv, bs, D, dtype=dtype)
            vc_h = torch.zeros_like(kc_h)
            # fill the CPU-layout cache THROUGH the reference's own cache-write kernel
            for s in range(S):
                L = lens[s]
                pos = torch.arange(L)
                slots = bt[s, pos // bs].long() * bs + pos % bs
                blocks = bt[s, pos // bs].long()
                ref_cpu.reshape_and_cache(kc[blocks, pos % bs].contiguous(), vc[blocks, pos % bs].contiguous(),
                                          kc_h, vc_h, slots, isa)
            case["out_ref_cpu_kernel"] = ref_cpu.decode_attention(q, kc_h, vc_h, bt, sl, scale, isa)
            case["cpu_isa"] = isa
        cases.append(case)
    torch.save(cases, OUT / "paged_attention.pt")

    # ---------------- cache write: the scatter loop of tests/kernels/attention/test_cache.py:400-414
    g = gen(7)
    cw = []
    for (T_, H, D, bs, nb, dtype) in [(42, 8, 128, 16, 4, torch.bfloat16), (7, 13, 46, 16, 2, torch.float16),
                                      (17, 2, 64, 32, 2, torch.float32)]:
        key = torch.randn(T_, H, D, generator=g).to(dtype)
        value = torch.randn(T_, H, D, generator=g).to(dtype)
        kc = torch.randn(nb, bs, H, D, generator=g).to(dtype)
        vc = torch.randn(nb, bs, H, D, generator=g).to(dtype)
        slots = torch.randperm(nb * bs, generator=g)[:T_].to(torch.int64)
        ek, ev = kc.clone(), vc.clone()
        block_indices = torch.div(slots, bs, rounding_mode="floor").tolist()
        block_offsets = (slots % bs).tolist()
        for i in range(T_):
            ek[block_indices[i], block_offsets[i], :, :] = key[i]
            ev[block_indices[i], block_offsets[i], :, :] = value[i]
        cw.append(dict(key=key, value=value, key_cache=kc, value_cache=vc, slot_mapping=slots,
                       expect_key_cache=ek, expect_value_cache=ev))
    torch.save(cw, OUT / "cache_write.pt")

    # ---------------- merge_attn_states
    ns = {"torch": torch}
    extract(T / "attention" / "test_merge_attn_states.py", ["merge_attn_states_torch"], ns)
    g = gen(11)
    mg = []
    for (Tn, H, D, dtype) in [(37, 4, 128, torch.bfloat16), (16, 8, 64, torch.float16)]:
        po = torch.randn(Tn, H, D, generator=g).to(dtype)
        so = torch.randn(Tn, H, D, generator=g).to(dtype)
        pl = torch.randn(H, Tn, generator=g) * 3
        sl_ = torch.randn(H, Tn, generator=g) * 3
        pl[0, 0] = math.inf   # inf handling
        sl_[1, 2] = -math.inf
        pl[2, 3] = -math.inf
        sl_[2, 3] = -math.inf  # both empty
        po[3, 2] = 0
        out, lse = ns["merge_attn_states_torch"](torch.empty_like(po), po.float(), pl.clone(), so.float(), sl_.clone(),
                                                 output_lse=torch.empty_like(pl))
        mg.append(dict(prefix_output=po, prefix_lse=pl, suffix_output=so, suffix_lse=sl_, out=out, out_lse=lse))
    torch.save(mg, OUT / "merge_attn_states.pt")

    # ---------------- FP8 quant + scaled_mm
    ns = {"torch": torch, "FP8_DTYPE": torch.float8_e4m3fn, "DEVICE": "cpu",
          "get_fp8_min_max": lambda: (-448.0, 448.0)}
    extract(T / "quant_utils.py", ["as_float32_tensor", "ref_dynamic_per_token_quant", "ref_dynamic_per_tensor_fp8_quant"], ns)
    extract(T / "utils.py", ["baseline_scaled_mm"], ns)
    g = gen(13)
    fq = []
    for (M, K, dtype, ub) in [(7, 4096, torch.bfloat16, None), (33, 1024, torch.float16, None),
                              (5, 17, torch.bfloat16, None), (9, 2560, torch.bfloat16, 2.0)]:
        x = (torch.randn(M, K, generator=g) * 3).to(dtype)
        x[0, :] = 0  # all-zero row -> min scaling factor
        ubt = None if ub is None else torch.tensor(ub, dtype=torch.float32)
        o_tok, s_tok = ns["ref_dynamic_per_token_quant"](x, torch.float8_e4m3fn, ubt)
        o_ten, s_ten = ns["ref_dynamic_per_tensor_fp8_quant"](x)
        fq.append(dict(x=x, scale_ub=ubt, per_token_out=o_tok.view(torch.uint8), per_token_scale=s_tok,
                       per_tensor_out=o_ten.view(torch.uint8), per_tensor_scale=s_ten))
    torch.save(fq, OUT / "fp8_quant.pt")
    mm = []
    for (M, N, K, per_tok, per_ch, use_bias) in [(1, 256, 128, True, True, False), (33, 512, 496, True, False, True),
                                                 (100, 1024, 256, False, True, False), (64, 128, 1024, True, True, True)]:
        a = (torch.randn(M, K, generator=g)).clamp(-448, 448).to(torch.float8_e4m3fn)
        b = (torch.randn(N, K, generator=g)).clamp(-448, 448).to(torch.float8_e4m3fn).t()  # [K,N] column-major
        sa = torch.rand((M, 1) if per_tok else (1, 1), generator=g) * 0.1 + 0.01
        sb = torch.rand((1, N) if per_ch else (1, 1), generator=g) * 0.1 + 0.01
        bias = (torch.rand(N, generator=g) * 10).to(torch.bfloat16) if use_bias else None
        out = ns["baseline_scaled_mm"](a, b, sa, sb, torch.bfloat16, bias)
        mm.append(dict(a=a.view(torch.uint8), b_nk=b.t().contiguous().view(torch.uint8), scale_a=sa, scale_b=sb,
                       bias=bias, out=out))
    torch.save(mm, OUT / "scaled_mm.pt")

    # ---------------- NVFP4 quant / swizzle / dequant-GEMM
    ns = {"torch": torch, "FLOAT4_E2M1_MAX": 6.0, "FLOAT8_E4M3_MAX": 448.0, "BLOCK_SIZE": 16}
    extract(T / "quantization" / "test_nvfp4_quant.py",
            ["cast_to_fp4", "get_reciprocal", "ref_nvfp4_quant", "recover_swizzled_scales", "E2M1_TO_FLOAT32", "cast_from_fp4"], ns)
    ns2 = {"torch": torch}
    extract(T / "quantization" / "nvfp4_utils.py",
            ["kE2M1ToFloat", "convert_swizzled_to_linear", "dequantize_nvfp4_to_dtype", "break_fp4_bytes"], ns2)
    ns3 = {"torch": torch, "round_up": lambda x, y: (x + y - 1) // y * y}
    src = (REF / "vllm/model_executor/layers/quantization/utils/nvfp4_utils.py").read_text()
    tree = ast.parse(src)
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "swizzle_blockscale"][0]
    code = ast.unparse(fn).replace(".contiguous().cuda()", ".contiguous()")  # CPU run: drop the device move only
    exec(code, ns3)
    g = gen(42)
    nq = []
    for (m, n, dtype) in [(128, 64, torch.float16), (150, 80, torch.bfloat16), (90, 48, torch.bfloat16),
                          (32, 1024, torch.bfloat16), (1, 64, torch.float16)]:
        x = torch.randn(m, n, generator=g).to(dtype)
        amax = x.abs().max().to(torch.float32)
        gs = (448.0 * 6.0 / amax).reshape(1)
        vals, sf = ns["ref_nvfp4_quant"](x, gs)
        sf_sw = ns3["swizzle_blockscale"](sf.to(torch.float8_e4m3fn))
        nq.append(dict(x=x, global_scale=gs, ref_values=vals, ref_scale=sf, ref_scale_swizzled=sf_sw.view(torch.uint8)))
    torch.save(nq, OUT / "nvfp4_quant.pt")
    fm = []
    for (m, n, k) in [(128, 128, 64), (150, 128, 64), (128, 160, 128), (16, 256, 512)]:
        a = torch.randn(m, k, generator=g).to(torch.bfloat16)
        b = torch.randn(n, k, generator=g).to(torch.bfloat16)
        gsa = (448.0 * 6.0 / a.abs().max().float()).reshape(1)
        gsb = (448.0 * 6.0 / b.abs().max().float()).reshape(1)
        # quantise with the reference's Python quantiser, pack like the kernel (low nibble = even index)
        def pack(x, gs):
            vals, sf = ns["ref_nvfp4_quant"](x, gs)
            mag = vals.abs()
            idx = torch.zeros_like(mag, dtype=torch.uint8)
            for i, v in enumerate([0.0, 0.5, 1.0, 1.5, 2.0, 3.0, 4.0, 6.0]):
                idx[mag == v] = i
            codes = idx | ((vals < 0) | ((vals == 0) & torch.signbit(vals))).to(torch.uint8) << 3
            packed = codes[:, 0::2] | (codes[:, 1::2] << 4)
            sw = ns3["swizzle_blockscale"](sf.to(torch.float8_e4m3fn))
            return packed, sw
        ap, asf = pack(a, gsa)
        bp, bsf = pack(b, gsb)
        a_dq = ns2["dequantize_nvfp4_to_dtype"](ap, asf, gsa, torch.float32, "cpu")
        b_dq = ns2["dequantize_nvfp4_to_dtype"](bp, bsf, gsb, torch.float32, "cpu")
        out = torch.matmul(a_dq, b_dq.t())
        fm.append(dict(a_packed=ap, b_packed=bp, sfa=asf.view(torch.uint8), sfb=bsf.view(torch.uint8),
                       alpha=(1.0 / (gsa * gsb)), out=out, a_dq=a_dq, b_dq=b_dq))
    torch.save(fm, OUT / "nvfp4_mm.pt")
    # ---------------- fused SiLU*mul + NVFP4 quant: the reference test's own recipe
    # (tests/kernels/quantization/test_silu_mul_nvfp4_quant.py:36-51: SiluAndMul.forward_native, then the NVFP4
    #  quantiser with global scale 448*6/amax; its bar is atol = rtol = 3e-1 on the dequantised values)
    import torch.nn.functional as F
    g = gen(42)
    sm = []
    for (m, n2, dtype) in [(128, 256, torch.float16), (128, 128, torch.bfloat16), (256, 256, torch.bfloat16),
                           (37, 96, torch.bfloat16)]:
        x = torch.randn(m, n2, generator=g).to(dtype)
        d = n2 // 2
        ref_out = F.silu(x[..., :d]) * x[..., d:]                     # SiluAndMul.forward_native (activation.py:134-137)
        gs = (448.0 * 6.0 / ref_out.abs().max().to(torch.float32)).reshape(1)
        vals, sf = ns["ref_nvfp4_quant"](ref_out, gs)
        sm.append(dict(x=x, global_scale=gs, ref_values=vals, ref_scale=sf))
    torch.save(sm, OUT / "silu_mul_nvfp4_quant.pt")
    total = sum(p.stat().st_size for p in OUT.glob("*.pt"))
    print(f"wrote {len(list(OUT.glob('*.pt')))} fixtures, {total / 1e6:.1f} MB, to {OUT}")


if __name__ == "__main__":
    main()
