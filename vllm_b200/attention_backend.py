"""vLLM attention backend over the sm_100a C-ABI kernels (drop-in boundary B1).

Registered as `AttentionBackendEnum.CUSTOM` (vllm/v1/attention/backends/registry.py:243-295)
by `vllm_b200.plugin.register()`; selected with `--attention-backend CUSTOM`
(`attention_config.backend`, vllm/v1/attention/selector.py:176-191).  The v1 model runner,
scheduler, KV-cache manager and `Attention` layer stay vLLM's: this file only supplies the
three classes the registry contract names (vllm/v1/attention/backend.py):

  AttentionBackend          :59-358    -> B200AttentionBackend
  AttentionMetadataBuilder  :584-746   -> B200AttentionMetadataBuilder   (reference impl
                                          vllm/v1/attention/backends/triton_attn.py:195-268)
  AttentionImpl             :869-980   -> B200AttentionImpl              (reference impls
                                          flash_attn.py:879-1176, triton_attn.py:541-795)

What runs where:
  do_kv_cache_update -> b200_reshape_and_cache_flash   (cache write over slot_mapping)
  forward            -> b200_paged_attention_decode    (split-KV + combine, GQA-grouped)

Pure-decode batches (max_query_len == 1) are the tuned case.  Batches that contain prefill /
chunked-prefill / spec-decode-verify tokens are served by the same kernel through the
`token_to_seq` indirection: every query token runs as its own causal single-query row with
length `num_computed + offset + 1`.  That is exact (causal attention of token i only sees
keys <= i, all of which are already in the paged cache because vLLM calls
`do_kv_cache_update` before `forward`, `forward_includes_kv_cache_update = False`), just not
tuned for long prefills.

Works against both the reference tree's API and the INSTALLED 0.22.0 wheel: the only
difference that matters is the KV-cache shape API, provided here for both.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import ClassVar

import torch
from vllm.v1.attention.backend import (
    AttentionBackend,
    AttentionCGSupport,
    AttentionImpl,
    AttentionMetadataBuilder,
    AttentionType,
    CommonAttentionMetadata,
    MultipleOf,
)

from . import ops

_FP8_KV = ("fp8", "fp8_e4m3", "fp8_e5m2")


def _kv_layout() -> str:
    """'NHD' (default) or 'HND' (vllm/v1/attention/backends/utils.py get_kv_cache_layout)."""
    try:
        from vllm.v1.attention.backends.utils import get_kv_cache_layout

        return get_kv_cache_layout()
    except Exception:
        return "NHD"


@dataclass
class B200AttentionMetadata:
    """Per-step metadata; every tensor is the runner's persistent buffer (CUDA-graph safe)."""

    num_actual_tokens: int
    max_query_len: int
    max_seq_len: int
    query_start_loc: torch.Tensor  # int32 [num_reqs + 1]
    seq_lens: torch.Tensor         # int32 [num_reqs]
    block_table: torch.Tensor      # int32 [num_reqs, max_blocks]
    slot_mapping: torch.Tensor     # int64 [num_tokens]
    # mixed batches only: one row per query token
    token_to_seq: torch.Tensor | None = None    # int32 [num_tokens] -> block_table row
    token_seq_lens: torch.Tensor | None = None  # int32 [num_tokens] causal length of each token
    # split-KV scratch owned by the builder (never allocated inside a capture)
    workspace: torch.Tensor | None = None
    num_splits: int = 0


class B200AttentionMetadataBuilder(AttentionMetadataBuilder[B200AttentionMetadata]):
    # decode-only and mixed batches both replay as graphs: the launch shape depends only on
    # (num_tokens, heads, max_seq_len bucket), never on device data
    _cudagraph_support: ClassVar[AttentionCGSupport] = AttentionCGSupport.ALWAYS
    reorder_batch_threshold: int | None = None

    def __init__(self, kv_cache_spec, layer_names, vllm_config, device):
        super().__init__(kv_cache_spec, layer_names, vllm_config, device)
        self.block_size = kv_cache_spec.block_size
        mc, pc = vllm_config.model_config, vllm_config.parallel_config
        self.num_heads_q = mc.get_num_attention_heads(pc)
        self.num_heads_kv = mc.get_num_kv_heads(pc)
        self.head_size = mc.get_head_size()
        self.max_model_len = mc.max_model_len
        spec_dtype = getattr(kv_cache_spec, "dtype", None)
        self._kv_kind = "fp8" if (spec_dtype is not None and torch.empty(0, dtype=spec_dtype).element_size() == 1) else "auto"
        sc = vllm_config.scheduler_config
        self.max_tokens = max(int(getattr(sc, "max_num_batched_tokens", 0) or 0), int(getattr(sc, "max_num_seqs", 0) or 0), 1)
        # persistent split-KV workspace sized for the worst (num_tokens, splits) the heuristic can pick
        self.workspace = torch.empty(self._max_workspace_bytes(), dtype=torch.uint8, device=device)
        # persistent per-token buffers for mixed batches (graph capture needs stable addresses)
        self._token_to_seq = torch.zeros(self.max_tokens, dtype=torch.int32, device=device)
        self._token_seq_lens = torch.zeros(self.max_tokens, dtype=torch.int32, device=device)
        self._arange = torch.arange(self.max_tokens, dtype=torch.int32, device=device)

    def _splits(self, num_tokens: int, max_seq_len: int) -> int:
        s = ops.paged_decode_num_splits(num_tokens, self.num_heads_q, self.num_heads_kv, max_seq_len, self.device,
                                        self.head_size, self._kv_kind)
        ws = getattr(self, "workspace", None)
        if ws is not None:  # never outgrow the persistent workspace (graph-captured address)
            while s > 1 and ops.paged_decode_workspace_bytes(num_tokens, self.num_heads_q, self.head_size, s) > ws.numel():
                s -= 1
        return s

    def _max_workspace_bytes(self) -> int:
        need, t = 16, 1
        while True:  # probe powers of two and their neighbours; `_splits` clamps to the buffer afterwards
            for tt in {t, min(self.max_tokens, 2 * t - 1), min(self.max_tokens, t + t // 2)}:
                for msl in {self.max_model_len, max(256, self.max_model_len // 4)}:
                    s = self._splits(tt, msl)
                    need = max(need, ops.paged_decode_workspace_bytes(tt, self.num_heads_q, self.head_size, s))
            if t >= self.max_tokens:
                break
            t = min(2 * t, self.max_tokens)
        return 2 * need

    def build(self, common_prefix_len: int, common_attn_metadata: CommonAttentionMetadata,
              fast_build: bool = False) -> B200AttentionMetadata:
        m = common_attn_metadata
        T = m.num_actual_tokens
        md = B200AttentionMetadata(
            num_actual_tokens=T, max_query_len=m.max_query_len, max_seq_len=m.max_seq_len,
            query_start_loc=m.query_start_loc, seq_lens=m.seq_lens, block_table=m.block_table_tensor,
            slot_mapping=m.slot_mapping, workspace=self.workspace)
        if m.max_query_len > 1:
            # token t of request r (query_start_loc[r] <= t < query_start_loc[r+1]) attends to
            # seq_lens[r] - (query_start_loc[r+1] - 1 - t) keys.  No host sync: searchsorted on device.
            qsl = m.query_start_loc
            tok = self._arange[:T]
            req = torch.searchsorted(qsl[1:].contiguous(), tok, right=True).clamp_(max=m.seq_lens.shape[0] - 1)
            self._token_to_seq[:T].copy_(req)
            lens = m.seq_lens[req] - (qsl[1:][req] - 1 - tok)
            # CUDA-graph pad tokens (tok >= query_start_loc[-1]) belong to no request: length 0, so the kernel
            # writes zeros for them and never walks block_table entries past the last request's blocks
            lens = torch.where(tok < qsl[-1], lens, torch.zeros_like(lens))
            self._token_seq_lens[:T].copy_(lens.clamp_(min=0))
            md.token_to_seq = self._token_to_seq[:T]
            md.token_seq_lens = self._token_seq_lens[:T]
        n_rows = T if m.max_query_len > 1 else m.seq_lens.shape[0]
        md.num_splits = self._splits(max(n_rows, 1), max(int(m.max_seq_len), 1))
        return md

    def build_for_cudagraph_capture(self, common_attn_metadata: CommonAttentionMetadata) -> B200AttentionMetadata:
        md = self.build(0, common_attn_metadata)
        # like triton_attn.py:185-193: capturing at max_model_len would walk huge block tables
        md.seq_lens.fill_(1)
        return md


class B200AttentionBackend(AttentionBackend):
    accept_output_buffer: bool = True
    supported_dtypes: ClassVar[list[torch.dtype]] = [torch.float16, torch.bfloat16]
    supported_kv_cache_dtypes: ClassVar[list[str]] = ["auto", "float16", "bfloat16", "fp8", "fp8_e4m3", "fp8_e5m2"]
    forward_includes_kv_cache_update: bool = False

    @staticmethod
    def get_name() -> str:
        return "CUSTOM"  # must be an AttentionBackendEnum member name (attention.py:423)

    @staticmethod
    def get_impl_cls():
        return B200AttentionImpl

    @staticmethod
    def get_builder_cls():
        return B200AttentionMetadataBuilder

    @staticmethod
    def get_supported_kernel_block_sizes():
        return [MultipleOf(16)]  # one TMA box = 16 tokens of one block

    @classmethod
    def get_supported_head_sizes(cls) -> list[int]:
        return [64, 128, 256]

    @classmethod
    def supports_compute_capability(cls, capability) -> bool:
        return capability.major == 10  # sm_100a only; there is no other code path

    @classmethod
    def supports_combination(cls, head_size, dtype, kv_cache_dtype, block_size, use_mla, has_sink, use_sparse,
                             device_capability) -> str | None:
        if head_size == 64 and kv_cache_dtype in _FP8_KV:
            return "head_size 64 with an fp8 KV cache is not built (row < 128 bytes)"
        return None

    # ---- KV-cache shape: INSTALLED 0.22.0 API (INSTALLED: v1/attention/backend.py:89-148)
    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int,
                           cache_dtype_str: str = "auto") -> tuple[int, ...]:
        if block_size % 16 != 0:
            raise ValueError("Block size must be a multiple of 16.")
        return (2, num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def get_kv_cache_stride_order(include_num_layers_dimension: bool = False) -> tuple[int, ...]:
        layout = _kv_layout()
        if include_num_layers_dimension:
            # logical (L, 2, B, N, H, D) -> physical (B, L, 2, N, H, D) / (B, L, 2, H, N, D)
            return (2, 0, 1, 3, 4, 5) if layout == "NHD" else (2, 0, 1, 4, 3, 5)
        return (0, 1, 2, 3, 4) if layout == "NHD" else (0, 1, 3, 2, 4)

    # ---- reference tree API (vllm/v1/attention/backend.py:350-354): both block layouts work,
    # the kernels take explicit (block, page, head) strides
    @classmethod
    def supported_kv_cache_layouts(cls):
        from vllm.v1.kv_cache_layout import KVCacheLayout  # only exists in the tree's version

        return (KVCacheLayout.LBNHC, KVCacheLayout.LBHNC)

    @staticmethod
    def use_cascade_attention(*args, **kwargs) -> bool:
        return False


class B200AttentionImpl(AttentionImpl[B200AttentionMetadata]):
    # The C ABI returns the softmax LSE (natural log, b200_paged_decode_params.out_lse; tested), but `forward` does not
    # run the decode-context-parallel flow (all-gather q -> local context attention -> cp_lse_ag_out_rs -> merge,
    # flash_attn.py:1178-1330), so the capability is NOT advertised: with DCP > 1 vLLM refuses this backend at
    # start-up (vllm/v1/attention/backend.py:791-816,861) instead of silently computing shard-local attention.
    can_return_lse_for_decode: bool = False
    supports_dcp: bool = False

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: int | None = None,
                 alibi_slopes: list[float] | None = None, sliding_window: int | None = None,
                 kv_cache_dtype: str = "auto", logits_soft_cap: float | None = None,
                 attn_type: str = AttentionType.DECODER, kv_sharing_target_layer_name: str | None = None,
                 **extra_impl_args) -> None:
        if alibi_slopes is not None:
            raise NotImplementedError("B200 attention backend: ALiBi is not built")
        if extra_impl_args.get("sinks") is not None:
            raise NotImplementedError("B200 attention backend: attention sinks are not built")
        if attn_type != AttentionType.DECODER:
            raise NotImplementedError("B200 attention backend: only decoder self-attention is built")
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        if self.num_heads % self.num_kv_heads:
            raise ValueError("num_heads must be a multiple of num_kv_heads")
        self.sliding_window = int(sliding_window) if sliding_window else 0
        self.kv_cache_dtype = kv_cache_dtype
        self.logits_soft_cap = float(logits_soft_cap or 0.0)
        self.attn_type = attn_type
        self.kv_sharing_target_layer_name = kv_sharing_target_layer_name
        self.supports_quant_query_input = False  # queries stay bf16/fp16
        ops.kv_dtype_code(kv_cache_dtype)  # raises on an unknown kv-cache dtype, like the reference

    # kv_cache: (2, num_blocks, block_size, H, D) logical view (possibly HND-strided)
    @staticmethod
    def _split_cache(kv_cache: torch.Tensor):
        if kv_cache.dim() != 5:
            raise RuntimeError(f"B200 attention backend: unexpected kv_cache rank {kv_cache.dim()}")
        if kv_cache.shape[0] == 2:
            return kv_cache[0], kv_cache[1]
        if kv_cache.shape[1] == 2:  # (num_blocks, 2, N, H, D) layouts (Triton backend shape)
            return kv_cache[:, 0], kv_cache[:, 1]
        raise RuntimeError(f"B200 attention backend: cannot find the K/V axis in kv_cache shape {tuple(kv_cache.shape)}")

    def do_kv_cache_update(self, layer, key: torch.Tensor, value: torch.Tensor, kv_cache: torch.Tensor,
                           slot_mapping: torch.Tensor) -> None:
        """Contract: vllm/v1/attention/backend.py:896-908; reference flash_attn.py:1142-1176."""
        if kv_cache.numel() == 0:
            return
        key_cache, value_cache = self._split_cache(kv_cache)
        # num_tokens = slot_mapping.size(0), key/value may be longer (CUDA-graph padding)
        ops.reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, self.kv_cache_dtype,
                                    layer._k_scale, layer._v_scale)

    # ---- RoPE + KV-cache write in one launch (vllm/v1/attention/backend.py:932-980).  Called by
    # torch.ops.vllm.fused_rope_and_unified_kv_cache_update (compilation/passes/fusion/rope_kvcache_fusion.py:39-70)
    # when RopeKVCacheFusionPass has replaced rotary_embedding + unified_kv_cache_update.  NOTE: the reference's own
    # config gate keeps that pass off outside ROCm (vllm/config/compilation.py:293-298), so on CUDA the hook is
    # reachable through the op / a direct call until that gate opens.
    def fused_rope_kvcache_supported(self) -> bool:
        return True

    def do_rope_and_kv_cache_update(self, layer, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                                    positions: torch.Tensor, cos_sin_cache: torch.Tensor, is_neox: bool,
                                    kv_cache: torch.Tensor, layer_slot_mapping: torch.Tensor) -> None:
        """In place: query/key rotated, (rotated key, value) written to the paged cache at layer_slot_mapping.
        Bit-identical to rotary_embedding followed by do_kv_cache_update (tests/test_gpu_vllm_plugin.py)."""
        if not is_neox or kv_cache.numel() == 0:
            # GPT-J style rotation is not built into the fused kernel: rotate with the reference's own op, then
            # our cache write (still one launch fewer than nothing, and always correct)
            from vllm import _custom_ops as vops

            vops.rotary_embedding(positions, query, key, self.head_size, cos_sin_cache, is_neox)
            if kv_cache.numel():
                self.do_kv_cache_update(layer, key.view(-1, self.num_kv_heads, self.head_size),
                                        value.view(-1, self.num_kv_heads, self.head_size), kv_cache, layer_slot_mapping)
            return
        key_cache, value_cache = self._split_cache(kv_cache)
        ops.rotary_embedding_and_cache_write(positions, query, key, value, self.head_size, cos_sin_cache, key_cache,
                                             value_cache, layer_slot_mapping, self.kv_cache_dtype, layer._k_scale,
                                             layer._v_scale)

    def forward(self, layer, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, kv_cache: torch.Tensor,
                attn_metadata: B200AttentionMetadata, output: torch.Tensor | None = None,
                output_scale: torch.Tensor | None = None, output_block_scale: torch.Tensor | None = None,
                ) -> torch.Tensor:
        """query [T, Hq, D] (or [T, Hq*D]); writes output[:num_actual_tokens] in place and returns it."""
        assert output is not None, "Output tensor must be provided."
        if output_block_scale is not None:
            raise NotImplementedError("B200 attention backend: fused NVFP4 (block-scaled) output quantisation is not built")
        if attn_metadata is None:  # profiling run (flash_attn.py:915-917)
            return output.fill_(0)
        md = attn_metadata
        T = md.num_actual_tokens
        q = query.view(-1, self.num_heads, self.head_size)[:T]
        o = output.view(-1, self.num_heads, self.head_size)[:T]
        key_cache, value_cache = self._split_cache(kv_cache)
        if md.token_to_seq is None:
            # pure decode: one row per request; T may be graph-padded beyond the request count
            n = min(T, md.seq_lens.shape[0])
            if n < T:
                o[n:].zero_()
                q, o = q[:n], o[:n]
            seq_lens, token_to_seq = md.seq_lens[:n], None
        else:
            seq_lens, token_to_seq = md.token_seq_lens, md.token_to_seq
        fp8 = self.kv_cache_dtype in _FP8_KV
        ops.paged_attention_decode(
            o, q, key_cache, value_cache, md.block_table, seq_lens, self.scale, md.max_seq_len, self.kv_cache_dtype,
            layer._k_scale if fp8 else None, layer._v_scale if fp8 else None, num_splits=md.num_splits,
            workspace=md.workspace, logits_soft_cap=self.logits_soft_cap, sliding_window=self.sliding_window,
            token_to_seq=token_to_seq, output_scale=output_scale)
        return output

    def fused_output_quant_supported(self, quant_key) -> bool:
        """AttnFusionPass asks before folding the static FP8 quant in front of o_proj into attention
        (vllm/v1/attention/backend.py:910-922; reference impls triton_attn.py:453, flashinfer.py:1349).  Static
        per-tensor symmetric FP8 is written straight from the split-combine / epilogue; NVFP4 output is not built."""
        try:
            from vllm.model_executor.layers.quantization.utils.quant_utils import kFp8StaticTensorSym
        except Exception:
            return False
        return quant_key == kFp8StaticTensorSym
