"""`torch.ops.b200.*`: the reference's custom-op schemas for the decode hot path, 1:1, bound
to the C ABI.

The reference registers these ops from C++ (`csrc/libtorch_stable/torch_bindings.cpp`); a
drop-in has to present the same argument lists so `vllm/_custom_ops.py`-style callers, fake
tensor tracing (torch.compile / CUDA-graph capture in vLLM's piecewise mode) and
`torch.library.opcheck` keep working.  Schemas mirrored (path:line under /root/reference):

  reshape_and_cache_flash                 torch_bindings.cpp:915-921
  cutlass_scaled_mm                       torch_bindings.cpp:116-119
  cutlass_scaled_fp4_mm                   torch_bindings.cpp:187-190
  scaled_fp4_quant.out                    torch_bindings.cpp:218-222
  static_scaled_fp8_quant                 torch_bindings.cpp:645-647
  dynamic_scaled_fp8_quant                torch_bindings.cpp:650-653
  dynamic_per_token_scaled_fp8_quant      torch_bindings.cpp:656-659
  merge_attn_states                       torch_bindings.cpp:354-362
  rms_norm / fused_add_rms_norm           torch_bindings.cpp:370-377
  rms_norm_dynamic_per_token_quant        torch_bindings.cpp:394-396
  rotary_embedding                        torch_bindings.cpp:418-421 (neox layout only)
  silu_and_mul                            torch_bindings.cpp:582
  silu_and_mul_nvfp4_quant                torch_bindings.cpp:251-253
  paged_attention_decode                  (ours; the legacy paged_attention_v1/v2 ops only exist
                                           in the installed wheel, INSTALLED:_custom_ops.py:114-205)

Only the CUDA dispatch key is implemented: on any other device the dispatcher raises, there
is no CPU fallback.  Fake ("meta") impls do nothing because every op writes into
caller-allocated outputs, exactly like the reference (`Tensor!` arguments, `-> ()`).
"""
from __future__ import annotations

import torch

from . import _lib, ops

NAMESPACE = "b200"
_lib_def = torch.library.Library(NAMESPACE, "FRAGMENT")

SCHEMAS = {
    "reshape_and_cache_flash":
        "(Tensor key, Tensor value, Tensor(a!) key_cache, Tensor(b!) value_cache, Tensor slot_mapping, "
        "str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()",
    "cutlass_scaled_mm":
        "(Tensor(a!) out, Tensor a, Tensor b, Tensor a_scales, Tensor b_scales, Tensor? bias) -> ()",
    "cutlass_scaled_fp4_mm":
        "(Tensor(a!) out, Tensor a, Tensor b, Tensor block_scale_a, Tensor block_scale_b, Tensor alpha) -> ()",
    "scaled_fp4_quant.out":
        "(Tensor input, Tensor input_scale, bool is_sf_swizzled_layout, *, Tensor(a!) output, "
        "Tensor(b!) output_scale) -> ()",
    "static_scaled_fp8_quant":
        "(Tensor(a!) result, Tensor input, Tensor scale, int[]? group_shape=None) -> ()",
    "dynamic_scaled_fp8_quant":
        "(Tensor(a!) result, Tensor input, Tensor(b!) scale) -> ()",
    "dynamic_per_token_scaled_fp8_quant":
        "(Tensor(a!) result, Tensor input, Tensor(b!) scale, Tensor? scale_ub) -> ()",
    "merge_attn_states":
        "(Tensor(a!) output, Tensor(b!)? output_lse, Tensor prefix_output, Tensor prefix_lse, "
        "Tensor suffix_output, Tensor suffix_lse, int? prefill_tokens_with_context=None, "
        "Tensor? output_scale=None) -> ()",
    "rms_norm": "(Tensor(a!) result, Tensor input, Tensor? weight, float epsilon) -> ()",
    "fused_add_rms_norm": "(Tensor(a!) input, Tensor(b!) residual, Tensor? weight, float epsilon) -> ()",
    "rms_norm_dynamic_per_token_quant":
        "(Tensor(a!) result, Tensor input, Tensor weight, Tensor(b!) scale, float epsilon, Tensor? scale_ub, "
        "Tensor(c!)? residual) -> ()",
    "rotary_embedding":
        "(Tensor positions, Tensor(a!) query, Tensor(b!)? key, int head_size, Tensor cos_sin_cache, bool is_neox, "
        "int rope_dim_offset=0, bool inverse=False) -> ()",
    "silu_and_mul": "(Tensor(a!) result, Tensor input) -> ()",
    "silu_and_mul_nvfp4_quant":
        "(Tensor(a!) result, Tensor(b!) result_block_scale, Tensor input, Tensor input_global_scale) -> ()",
    "paged_attention_decode":
        "(Tensor(a!) out, Tensor query, Tensor key_cache, Tensor value_cache, Tensor block_table, Tensor seq_lens, "
        "float scale, int max_seq_len, str kv_cache_dtype, Tensor? k_scale, Tensor? v_scale, int num_splits, "
        "Tensor(b!)? workspace, Tensor(c!)? out_lse, float logits_soft_cap, int sliding_window, "
        "Tensor? token_to_seq) -> ()",
}


# ------------------------------------------------------------------------------ CUDA impls
def _reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale):
    ops.reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale)


def _cutlass_scaled_mm(out, a, b, a_scales, b_scales, bias):
    ops.cutlass_scaled_mm(a, b, a_scales, b_scales, out.dtype, bias, out=out)


def _cutlass_scaled_fp4_mm(out, a, b, block_scale_a, block_scale_b, alpha):
    ops.cutlass_scaled_fp4_mm(a, b, block_scale_a, block_scale_b, alpha, out.dtype, out=out)


def _scaled_fp4_quant_out(input, input_scale, is_sf_swizzled_layout, *, output, output_scale):
    ops.scaled_fp4_quant_out(input, input_scale, is_sf_swizzled_layout, output, output_scale)


def _static_scaled_fp8_quant(result, input, scale, group_shape=None):
    ops.scaled_fp8_quant(input, scale, output=result, group_shape=tuple(group_shape) if group_shape else None)


def _dynamic_scaled_fp8_quant(result, input, scale):
    ops.scaled_fp8_quant(input, None, output=result, scale_out=scale)


def _dynamic_per_token_scaled_fp8_quant(result, input, scale, scale_ub):
    ops.scaled_fp8_quant(input, None, scale_ub=scale_ub, use_per_token_if_dynamic=True, output=result, scale_out=scale)


def _merge_attn_states(output, output_lse, prefix_output, prefix_lse, suffix_output, suffix_lse,
                       prefill_tokens_with_context=None, output_scale=None):
    if output_scale is not None:
        raise RuntimeError("b200::merge_attn_states: fused fp8 output quantisation (output_scale) is not built")
    if prefill_tokens_with_context is not None and prefill_tokens_with_context < output.shape[0]:
        # tokens >= prefill_tokens_with_context have no prefix: output = suffix (merge_attn_states.cu:47-60)
        n = int(prefill_tokens_with_context)
        output[n:].copy_(suffix_output[n:])
        if output_lse is not None:
            output_lse[:, n:].copy_(suffix_lse[:, n:])
        if n == 0:
            return
        ops.merge_attn_states(output[:n], prefix_output[:n], prefix_lse[:, :n].contiguous(), suffix_output[:n],
                              suffix_lse[:, :n].contiguous(), None)
        if output_lse is not None:
            tmp = torch.empty_like(prefix_lse[:, :n].contiguous())
            ops.merge_attn_states(output[:n], prefix_output[:n], prefix_lse[:, :n].contiguous(), suffix_output[:n],
                                  suffix_lse[:, :n].contiguous(), tmp)
            output_lse[:, :n].copy_(tmp)
        return
    ops.merge_attn_states(output, prefix_output, prefix_lse, suffix_output, suffix_lse, output_lse)


def _paged_attention_decode(out, query, key_cache, value_cache, block_table, seq_lens, scale, max_seq_len,
                            kv_cache_dtype, k_scale, v_scale, num_splits, workspace, out_lse, logits_soft_cap,
                            sliding_window, token_to_seq):
    ops.paged_attention_decode(out, query, key_cache, value_cache, block_table, seq_lens, scale, max_seq_len,
                               kv_cache_dtype, k_scale, v_scale, num_splits, workspace, out_lse, logits_soft_cap,
                               sliding_window, token_to_seq)


def _need_weight(weight, what):
    if weight is None:
        raise RuntimeError(f"b200::{what}: weight=None (norm without scale) is not built")
    return weight


def _rms_norm(result, input, weight, epsilon):
    ops.rms_norm(result, input, _need_weight(weight, "rms_norm"), epsilon)


def _fused_add_rms_norm(input, residual, weight, epsilon):
    ops.fused_add_rms_norm(input, residual, _need_weight(weight, "fused_add_rms_norm"), epsilon)


def _rms_norm_dynamic_per_token_quant(result, input, weight, scale, epsilon, scale_ub, residual):
    ops.rms_norm_dynamic_per_token_quant(input, weight, epsilon, scale_ub=scale_ub, residual=residual, output=result,
                                         scales=scale)


def _rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox, rope_dim_offset=0, inverse=False):
    if rope_dim_offset != 0 or inverse or key is None:
        raise RuntimeError("b200::rotary_embedding: rope_dim_offset / inverse / key=None are not built")
    ops.rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox)


def _silu_and_mul(result, input):
    ops.silu_and_mul(result, input)


def _silu_and_mul_nvfp4_quant(result, result_block_scale, input, input_global_scale):
    ops.silu_and_mul_nvfp4_quant(result, result_block_scale, input, input_global_scale)


_IMPLS = {
    "rms_norm": _rms_norm,
    "fused_add_rms_norm": _fused_add_rms_norm,
    "rms_norm_dynamic_per_token_quant": _rms_norm_dynamic_per_token_quant,
    "rotary_embedding": _rotary_embedding,
    "silu_and_mul": _silu_and_mul,
    "silu_and_mul_nvfp4_quant": _silu_and_mul_nvfp4_quant,
    "reshape_and_cache_flash": _reshape_and_cache_flash,
    "cutlass_scaled_mm": _cutlass_scaled_mm,
    "cutlass_scaled_fp4_mm": _cutlass_scaled_fp4_mm,
    "scaled_fp4_quant.out": _scaled_fp4_quant_out,
    "static_scaled_fp8_quant": _static_scaled_fp8_quant,
    "dynamic_scaled_fp8_quant": _dynamic_scaled_fp8_quant,
    "dynamic_per_token_scaled_fp8_quant": _dynamic_per_token_scaled_fp8_quant,
    "merge_attn_states": _merge_attn_states,
    "paged_attention_decode": _paged_attention_decode,
}


def _noop_fake(*args, **kwargs):
    return None


_registered = False


def register() -> None:
    """Define the ops once per process (idempotent)."""
    global _registered
    if _registered:
        return
    for name, schema in SCHEMAS.items():
        _lib_def.define(name + schema)
        _lib_def.impl(name, _IMPLS[name], "CUDA")
        torch.library.register_fake(f"{NAMESPACE}::{name}", _noop_fake, lib=_lib_def)
    _registered = True


register()
