"""Drop-in boundary checks that need no GPU: the plugin registers through vLLM's own
registries, the backend answers the selector's questions like a vLLM backend, the torch.ops
schemas equal the reference's, and the metadata builder's mixed-batch index math is exact."""
import re
import types

import pytest
import torch

vllm = pytest.importorskip("vllm")


@pytest.fixture(scope="module")
def registered():
    from vllm_b200 import plugin

    plugin.register()
    plugin.register()  # idempotent: vLLM may load general plugins more than once per process
    return plugin


def test_entry_point_is_discoverable():
    from importlib.metadata import entry_points

    eps = {e.name: e.value for e in entry_points(group="vllm.general_plugins")}
    assert eps.get("b200") == "vllm_b200.plugin:register"


def test_backend_registered_as_custom(registered):
    from vllm.v1.attention.backends.registry import AttentionBackendEnum

    assert AttentionBackendEnum.CUSTOM.get_path() == registered.BACKEND_PATH
    cls = AttentionBackendEnum.CUSTOM.get_class()
    assert cls.get_name() == "CUSTOM" and cls.get_name() in AttentionBackendEnum.__members__
    assert cls.forward_includes_kv_cache_update is False
    assert cls.get_impl_cls().__name__ == "B200AttentionImpl"
    assert cls.get_builder_cls().__name__ == "B200AttentionMetadataBuilder"
    assert cls.get_kv_cache_shape(10, 16, 8, 128, "fp8") == (2, 10, 16, 8, 128)
    assert cls.get_kv_cache_block_dim(16, 8, 128) == 1
    with pytest.raises(ValueError):
        cls.get_kv_cache_shape(10, 8, 8, 128)
    assert sorted(cls.get_kv_cache_stride_order()) == [0, 1, 2, 3, 4]
    assert sorted(cls.get_kv_cache_stride_order(include_num_layers_dimension=True)) == list(range(6))


def test_backend_validate_configuration(registered):
    from vllm.platforms.interface import DeviceCapability
    from vllm.v1.attention.backends.registry import AttentionBackendEnum

    cls = AttentionBackendEnum.CUSTOM.get_class()
    base = dict(head_size=128, dtype=torch.bfloat16, kv_cache_dtype="fp8", block_size=16, use_mla=False,
                has_sink=False, use_sparse=False, use_mm_prefix=False, use_per_head_quant_scales=False,
                device_capability=DeviceCapability(10, 0), attn_type="decoder")
    assert cls.validate_configuration(**base) == []
    assert cls.validate_configuration(**{**base, "device_capability": DeviceCapability(9, 0)}) == [
        "compute capability not supported"]
    assert "head_size not supported" in cls.validate_configuration(**{**base, "head_size": 96})
    assert "block_size not supported" in cls.validate_configuration(**{**base, "block_size": 8})
    assert "dtype not supported" in cls.validate_configuration(**{**base, "dtype": torch.float32})
    assert "MLA not supported" in cls.validate_configuration(**{**base, "use_mla": True})
    assert "attention sinks not supported" in cls.validate_configuration(**{**base, "has_sink": True})
    assert cls.validate_configuration(**{**base, "head_size": 64}) != []       # fp8 rows < 128 B
    assert cls.validate_configuration(**{**base, "head_size": 64, "kv_cache_dtype": "auto"}) == []


def test_linear_kernels_first_in_vllm_tables(registered):
    from vllm.model_executor.kernels import linear as L
    from vllm.platforms.interface import PlatformEnum

    assert L._POSSIBLE_FP8_KERNELS[PlatformEnum.CUDA][0].__name__ == "B200FP8ScaledMMLinearKernel"
    assert L._POSSIBLE_NVFP4_KERNELS[PlatformEnum.CUDA][0].__name__ == "B200NvFp4LinearKernel"
    # without an sm_100 device the classes must say "not supported" (never silently run elsewhere)
    ok, why = L._POSSIBLE_FP8_KERNELS[PlatformEnum.CUDA][0].is_supported(90)
    assert not ok and why
    if not torch.cuda.is_available():
        ok, why = L._POSSIBLE_NVFP4_KERNELS[PlatformEnum.CUDA][0].is_supported()
        assert not ok and "CUDA" in why


# the reference's schemas, as written at csrc/libtorch_stable/torch_bindings.cpp (line numbers in torch_ops.py)
REFERENCE_SCHEMAS = {
    "reshape_and_cache_flash": "reshape_and_cache_flash(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache, "
                               "Tensor slot_mapping, str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()",
    "cutlass_scaled_mm": "cutlass_scaled_mm(Tensor! out, Tensor a, Tensor b, Tensor a_scales, Tensor b_scales, Tensor? bias) -> ()",
    "cutlass_scaled_fp4_mm": "cutlass_scaled_fp4_mm(Tensor! out, Tensor a, Tensor b, Tensor block_scale_a, "
                             "Tensor block_scale_b, Tensor alpha) -> ()",
    "scaled_fp4_quant.out": "scaled_fp4_quant.out(Tensor input, Tensor input_scale, bool is_sf_swizzled_layout, *, "
                            "Tensor(a!) output, Tensor(b!) output_scale) -> ()",
    "static_scaled_fp8_quant": "static_scaled_fp8_quant(Tensor! result, Tensor input, Tensor scale, int[]? group_shape=None) -> ()",
    "dynamic_scaled_fp8_quant": "dynamic_scaled_fp8_quant(Tensor! result, Tensor input, Tensor! scale) -> ()",
    "dynamic_per_token_scaled_fp8_quant": "dynamic_per_token_scaled_fp8_quant(Tensor! result, Tensor input, Tensor! scale, "
                                          "Tensor? scale_ub) -> ()",
    # torch_bindings.cpp:370-377, 394-396, 418-421, 582, 251-253
    "rms_norm": "rms_norm(Tensor! result, Tensor input, Tensor? weight, float epsilon) -> ()",
    "fused_add_rms_norm": "fused_add_rms_norm(Tensor! input, Tensor! residual, Tensor? weight, float epsilon) -> ()",
    "rms_norm_dynamic_per_token_quant": "rms_norm_dynamic_per_token_quant(Tensor! result, Tensor input, Tensor weight, "
                                        "Tensor! scale, float epsilon, Tensor? scale_ub, Tensor!? residual) -> ()",
    "rotary_embedding": "rotary_embedding(Tensor positions, Tensor! query, Tensor!? key, int head_size, "
                        "Tensor cos_sin_cache, bool is_neox, int rope_dim_offset=0, bool inverse=False) -> ()",
    "silu_and_mul": "silu_and_mul(Tensor! result, Tensor input) -> ()",
    "silu_and_mul_nvfp4_quant": "silu_and_mul_nvfp4_quant(Tensor! result, Tensor! result_block_scale, Tensor input, "
                                "Tensor input_global_scale) -> ()",
}


def _args(schema: str):
    """[(type-without-alias-annotation, name, is_mutable)] of a schema string."""
    inner = schema[schema.index("(") + 1: schema.rindex(") ->")]
    out = []
    for part in re.split(r",\s*(?![^()]*\))", inner):
        part = part.strip()
        if part == "*":
            continue
        typ, name = part.rsplit(" ", 1)
        mut = "!" in typ
        typ = re.sub(r"\([a-z]!\)|!", "", typ)
        out.append((typ, name.split("=")[0], mut))
    return out


def test_torch_ops_schemas_match_reference(registered):
    for name, ref in REFERENCE_SCHEMAS.items():
        base, _, overload = name.partition(".")
        op = getattr(getattr(torch.ops.b200, base), overload or "default")
        assert _args(str(op._schema)) == _args(ref), name
    s = str(torch.ops.b200.merge_attn_states.default._schema)
    assert [a[1] for a in _args(s)][:6] == ["output", "output_lse", "prefix_output", "prefix_lse", "suffix_output",
                                             "suffix_lse"]


def test_torch_ops_have_no_cpu_kernel(registered):
    out = torch.empty(2, 16, dtype=torch.bfloat16)
    a = torch.zeros(2, 16, dtype=torch.float8_e4m3fn)
    with pytest.raises((NotImplementedError, RuntimeError)):
        torch.ops.b200.cutlass_scaled_mm(out, a, a.t(), torch.ones(1), torch.ones(1), None)


def _fake_config(hq=8, hkv=2, d=128, max_len=4096, max_tokens=64):
    mc = types.SimpleNamespace(max_model_len=max_len, get_num_attention_heads=lambda pc: hq,
                               get_num_kv_heads=lambda pc: hkv, get_head_size=lambda: d)
    sc = types.SimpleNamespace(max_num_batched_tokens=max_tokens, max_num_seqs=16)
    return types.SimpleNamespace(model_config=mc, parallel_config=None, scheduler_config=sc)


def test_metadata_builder_mixed_batch_index_math(registered):
    """token -> (block_table row, causal length) for a prefill + decode + chunked-prefill batch."""
    from vllm.v1.attention.backend import CommonAttentionMetadata
    from vllm_b200.attention_backend import B200AttentionMetadataBuilder

    spec = types.SimpleNamespace(block_size=16)
    b = B200AttentionMetadataBuilder(spec, ["l0"], _fake_config(), torch.device("cpu"))
    # req0: decode (1 new token, 40 total); req1: prefill of 5 tokens (5 total); req2: chunk of 3 with 10 computed
    qsl = torch.tensor([0, 1, 6, 9], dtype=torch.int32)
    seq_lens = torch.tensor([40, 5, 13], dtype=torch.int32)
    cm = CommonAttentionMetadata(
        query_start_loc=qsl, query_start_loc_cpu=qsl, seq_lens=seq_lens, num_reqs=3, num_actual_tokens=9,
        max_query_len=5, max_seq_len=40, block_table_tensor=torch.zeros(3, 4, dtype=torch.int32),
        slot_mapping=torch.zeros(9, dtype=torch.int64))
    md = b.build(0, cm)
    assert md.token_to_seq.tolist() == [0, 1, 1, 1, 1, 1, 2, 2, 2]
    assert md.token_seq_lens.tolist() == [40, 1, 2, 3, 4, 5, 11, 12, 13]
    assert md.num_splits >= 1 and md.workspace.numel() >= 16
    # pure decode: no indirection, seq_lens used as is
    qsl = torch.arange(4, dtype=torch.int32)
    cm = cm.replace(query_start_loc=qsl, query_start_loc_cpu=qsl, num_actual_tokens=3, max_query_len=1,
                    slot_mapping=torch.zeros(3, dtype=torch.int64))
    md = b.build(0, cm)
    assert md.token_to_seq is None and md.seq_lens is seq_lens


def test_impl_rejects_what_is_not_built(registered):
    from vllm_b200.attention_backend import B200AttentionImpl

    with pytest.raises(NotImplementedError):
        B200AttentionImpl(8, 128, 0.1, 2, alibi_slopes=[1.0] * 8)
    with pytest.raises(NotImplementedError):
        B200AttentionImpl(8, 128, 0.1, 2, sinks=torch.zeros(8))
    with pytest.raises(RuntimeError):
        B200AttentionImpl(8, 128, 0.1, 2, kv_cache_dtype="int3")
    impl = B200AttentionImpl(8, 128, 0.1, 2, None, None, "fp8")
    out = torch.ones(4, 8 * 128)
    assert impl.forward(None, out, out, out, torch.empty(0), None, output=out).abs().sum() == 0  # profiling run
