"""vLLM linear-kernel classes over the sm_100a tcgen05 GEMMs (drop-in boundary B2).

vLLM's quantisation LinearMethods (`Fp8LinearMethod` layers/quantization/fp8.py:239-461,
`Fp8PerTensorOnlineLinearMethod` online/fp8.py:158-257, `ModelOptNvFp4LinearMethod`
modelopt.py:1109-1240) already delegate `process_weights_after_loading` / `apply` to a kernel
object picked from `_POSSIBLE_FP8_KERNELS` / `_POSSIBLE_NVFP4_KERNELS`
(vllm/model_executor/kernels/linear/__init__.py:397-428,526-543).  These two classes are that
kernel object; `plugin.register()` puts them FIRST for PlatformEnum.CUDA
(`register_linear_kernel` only appends = lowest priority, :1110-1159).

  B200FP8ScaledMMLinearKernel   replaces CutlassFP8ScaledMMLinearKernel (scaled_mm/cutlass.py:156-272)
  B200NvFp4LinearKernel         replaces CutlassNvFp4LinearKernel      (nvfp4/cutlass.py:20-77)
"""
from __future__ import annotations

import os
from collections.abc import Sequence

import torch
from vllm.model_executor.kernels.linear import (
    FP8ScaledMMLinearKernel,
    FP8ScaledMMLinearLayerConfig,
    NvFp4LinearKernel,
    NvFp4LinearLayerConfig,
)

from . import ops
from . import torch_ops  # noqa: F401  (defines torch.ops.b200.*)

# vLLM traces the model forward with torch.compile (Dynamo): raw ctypes calls are not traceable, so the
# kernel classes go through the torch.ops.b200.* custom ops (opaque, mutation-annotated, with fake impls),
# exactly like the reference's classes go through torch.ops._C.*.
_B = torch.ops.b200
_OWN_QUANT = os.environ.get("B200_VLLM_QUANT", "vllm") == "b200"


def _is_sm100(compute_capability: int | None) -> tuple[bool, str | None]:
    if not torch.cuda.is_available():
        return False, "requires a CUDA device (sm_100a)"
    if compute_capability is None:
        major, minor = torch.cuda.get_device_capability()
        compute_capability = major * 10 + minor
    if compute_capability // 10 != 10:
        return False, f"built for sm_100a only, device is sm_{compute_capability}"
    try:
        from . import _lib

        _lib.load()
    except Exception as e:  # library missing: never pretend to be available
        return False, str(e)
    return True, None


class B200FP8ScaledMMLinearKernel(FP8ScaledMMLinearKernel):
    """FP8 W8A8: per-tensor or per-channel weight scales, static per-tensor or dynamic
    per-token / per-tensor activation scales (the combinations cutlass_scaled_mm accepts,
    scaled_mm_entry.cu:203-216)."""

    def __init__(self, c: FP8ScaledMMLinearLayerConfig, layer_param_names: Sequence[str]) -> None:
        self.logical_output_size: int | None = None
        super().__init__(c, layer_param_names)
        d = c.activation_quant_key.scale
        self._act_static = bool(d.static)
        gs = d.group_shape
        # GroupShape(1, -1) = per token; (-1, -1) = per tensor
        self._act_per_token = (not self._act_static) and tuple(gs)[0] == 1

    @classmethod
    def is_supported(cls, compute_capability: int | None = None) -> tuple[bool, str | None]:
        return _is_sm100(compute_capability)

    @classmethod
    def can_implement(cls, c: FP8ScaledMMLinearLayerConfig) -> tuple[bool, str | None]:
        for key, what in ((c.weight_quant_key, "weight"), (c.activation_quant_key, "activation")):
            gs = tuple(key.scale.group_shape)
            if gs not in ((-1, -1), (1, -1), (-1, 1)):
                return False, f"block/group-wise {what} scales {gs} are not built (per-tensor / per-token / per-channel only)"
        if c.out_dtype not in (None, torch.bfloat16, torch.float16):
            return False, f"out_dtype {c.out_dtype} not built"
        return True, None

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        w_name = self.layer_param_names[0]
        w = getattr(layer, w_name)  # [K, N] with stride(0) == 1 (fp8.py:405-409)
        self.logical_output_size = w.shape[1]
        if w.shape[0] % 16 or w.shape[1] % 16:
            raise ValueError(f"B200 FP8 GEMM needs K and N to be multiples of 16, got {tuple(w.shape)}")

    def apply_weights(self, layer: torch.nn.Module, x: torch.Tensor, bias: torch.Tensor | None = None) -> torch.Tensor:
        """Default: the base class's path, exactly what CutlassFP8ScaledMMLinearKernel does (ScaledMMLinearKernel.py
        :120-160) - the activation quant stays vLLM's `QuantFP8`, which under torch.compile is plain torch ops that
        Inductor FUSES into the preceding RMSNorm / SiLU kernel; only the GEMM is replaced.  Measured inside vLLM on
        B200 (profiles/r02_vllm_onoff.md): quantising through our opaque custom op instead cost 4 extra launches of
        7-10 us per layer.  B200_VLLM_QUANT=b200 selects our own quant kernels (torch.ops.b200.*_scaled_fp8_quant)."""
        if not _OWN_QUANT:
            return super().apply_weights(layer, x, bias)
        w, w_s, x_s, x_s_ub = self._get_layer_params(layer)
        x2 = x.reshape(-1, x.shape[-1])
        out_dtype = self.config.out_dtype or x.dtype
        if x2.dtype != self.fp8_dtype:
            xq = torch.empty(x2.shape, dtype=self.fp8_dtype, device=x2.device)
            if self._act_static:
                xs = x_s
                _B.static_scaled_fp8_quant(xq, x2, xs)
            elif self._act_per_token:
                xs = torch.empty((x2.shape[0], 1), dtype=torch.float32, device=x2.device)
                _B.dynamic_per_token_scaled_fp8_quant(xq, x2, xs, x_s_ub)
            else:
                xs = torch.empty(1, dtype=torch.float32, device=x2.device)
                _B.dynamic_scaled_fp8_quant(xq, x2, xs)
        else:
            xq, xs = x2, x_s
        return self.apply_scaled_mm(A=xq, B=w, out_dtype=out_dtype, As=xs, Bs=w_s, bias=bias,
                                    output_shape=[*x.shape[:-1], w.shape[1]])

    def apply_scaled_mm(self, *, A: torch.Tensor, B: torch.Tensor, out_dtype: torch.dtype, As: torch.Tensor,
                        Bs: torch.Tensor, bias: torch.Tensor | None, output_shape: list) -> torch.Tensor:
        out = torch.empty((A.shape[0], B.shape[1]), dtype=out_dtype, device=A.device)
        _B.cutlass_scaled_mm(out, A, B, As.reshape(-1), Bs.reshape(-1), bias)
        return out.view(*output_shape)


class B200NvFp4LinearKernel(NvFp4LinearKernel):
    """NVFP4 W4A4 (ModelOpt layout): weight u8 [N, K/2], weight_scale e4m3 [N, K/16]
    linear -> swizzled once here, alpha = input_scale * weight_scale_2."""

    @classmethod
    def is_supported(cls, compute_capability: int | None = None) -> tuple[bool, str | None]:
        return _is_sm100(compute_capability)

    @classmethod
    def can_implement(cls, config: NvFp4LinearLayerConfig) -> tuple[bool, str | None]:
        return True, None

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        n, k2 = layer.weight.shape
        if n % 32 or (2 * k2) % 32:
            raise ValueError(f"B200 NVFP4 GEMM needs N and K to be multiples of 32, got N={n}, K={2 * k2}")
        layer.weight_scale = torch.nn.Parameter(ops.swizzle_blockscale(layer.weight_scale.data), requires_grad=False)
        layer.weight = torch.nn.Parameter(layer.weight.data.contiguous(), requires_grad=False)

    def apply_weights(self, layer: torch.nn.Module, x: torch.Tensor, bias: torch.Tensor | None = None) -> torch.Tensor:
        n = layer.output_size_per_partition
        x2 = x.reshape(-1, x.shape[-1])
        m, k = x2.shape
        xq = torch.empty((m, k // 2), dtype=torch.uint8, device=x2.device)
        # swizzled SF buffer [ru(m,128), ru(k/16,4)] e4m3; the quant kernel writes every padded entry
        xsf = torch.empty((ops.round_up(m, 128), ops.round_up(k // 16, 4)), dtype=torch.uint8, device=x2.device)
        _B.scaled_fp4_quant.out(x2, layer.input_global_scale_inv, True, output=xq, output_scale=xsf)
        out = torch.empty((m, layer.weight.shape[0]), dtype=x.dtype, device=x2.device)
        _B.cutlass_scaled_fp4_mm(out, xq, layer.weight, xsf, layer.weight_scale, layer.alpha)
        if out.shape[1] != n:
            out = out[:, :n].contiguous()
        if bias is not None:
            out = out + bias
        return out.view(*x.shape[:-1], n)
