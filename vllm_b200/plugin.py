"""`vllm.general_plugins` entry point: makes the B200 decode hot path selectable in an
UNMODIFIED vLLM (reference tree or the installed 0.22.0 wheel).

vLLM calls every `vllm.general_plugins` entry point in every process it starts (engine core,
each TP worker) through `load_general_plugins()` (vllm/plugins/__init__.py:36-90), so the
registrations below exist wherever a model is built.  Idempotent by design (the loader may
run more than once per process).

  1. attention backend   register_backend(AttentionBackendEnum.CUSTOM, "...B200AttentionBackend")
                         (vllm/v1/attention/backends/registry.py:243-295); used when the user passes
                         `--attention-backend CUSTOM` / `attention_config={"backend": "CUSTOM"}`.
  2. linear kernels      B200 FP8 / NVFP4 kernel classes inserted at index 0 of
                         `_POSSIBLE_FP8_KERNELS[CUDA]` / `_POSSIBLE_NVFP4_KERNELS[CUDA]`
                         (vllm/model_executor/kernels/linear/__init__.py:397-428,526-543) so that vLLM's own
                         Fp8LinearMethod / ModelOptNvFp4LinearMethod pick them.  Opt out per family with
                         B200_VLLM_LINEAR=0 or VLLM_DISABLED_KERNELS=B200FP8ScaledMMLinearKernel,...
  3. torch.ops.b200.*    the reference's op schemas bound to the C ABI (vllm_b200/torch_ops.py).
"""
from __future__ import annotations

import logging
import os

logger = logging.getLogger("vllm_b200")

BACKEND_PATH = "vllm_b200.attention_backend.B200AttentionBackend"
_done = False


def register() -> None:
    global _done
    if _done:
        return
    _done = True
    from . import torch_ops  # noqa: F401  (defines torch.ops.b200.*)

    from vllm.v1.attention.backends.registry import AttentionBackendEnum, register_backend

    register_backend(AttentionBackendEnum.CUSTOM, BACKEND_PATH)

    if os.environ.get("B200_VLLM_LINEAR", "1") != "0":
        from vllm.model_executor.kernels import linear as L
        from vllm.platforms.interface import PlatformEnum

        from .linear_kernels import B200FP8ScaledMMLinearKernel, B200NvFp4LinearKernel

        for table, cls in ((L._POSSIBLE_FP8_KERNELS, B200FP8ScaledMMLinearKernel),
                           (L._POSSIBLE_NVFP4_KERNELS, B200NvFp4LinearKernel)):
            lst = table.setdefault(PlatformEnum.CUDA, [])
            if cls not in lst:
                lst.insert(0, cls)  # highest priority; is_supported() keeps it off non-sm_100 devices
    logger.info("vllm_b200 plugin registered: attention backend CUSTOM -> %s; FP8/NVFP4 linear kernels first", BACKEND_PATH)
