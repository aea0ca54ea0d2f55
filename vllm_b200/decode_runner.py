"""Minimal tensor-parallel decode-step runner over the C-ABI kernels.

This is NOT vLLM's model runner (that stays vLLM's — see INTEGRATION.md); it is the
smallest harness that executes exactly the decode step of SURVEY §3.1 for a Llama-style
dense model so bench.py / smoke() / tests can time and check the hot path end to end:

  per layer:  [fused-add RMSNorm + FP8 per-token quant] -> qkv GEMM -> [RoPE + KV-cache write
              (slot_mapping)] -> paged-attention decode (block_table) -> quant -> o_proj GEMM ->
              [all-reduce] -> [fused-add RMSNorm + quant] -> gate_up GEMM -> [SiLU*mul + quant] ->
              down GEMM -> [all-reduce]          ([..] = one fused launch; fused=False runs them apart)
  then:       final norm -> lm_head (vocab-parallel library GEMM) -> greedy argmax.

Tensor parallelism follows vLLM's sharding (vllm/model_executor/layers/linear.py: QKV /
gate_up column-parallel by head / column, o_proj / down_proj row-parallel + all-reduce at
linear.py:1657-1660): one process per GPU, `torch.distributed` NCCL all-reduce, every rank
owns the KV cache of its own kv-heads with the same block_table / slot_mapping.
Weights are synthetic ("dummy" load format) and quantised online per tensor like
Fp8PerTensorOnlineLinearMethod (vllm/model_executor/layers/quantization/online/fp8.py:158-224).
"""
from __future__ import annotations

import dataclasses
import math
import os
import sys

import torch
import torch.distributed as dist

from . import _lib, ops, tp_shard


@dataclasses.dataclass
class ModelShape:
    name: str = "llama-3.1-8b"
    num_layers: int = 32
    hidden: int = 4096
    num_q_heads: int = 32
    num_kv_heads: int = 8
    head_size: int = 128
    intermediate: int = 14336
    vocab: int = 128256
    rope_theta: float = 500000.0
    rms_eps: float = 1e-5


LLAMA_8B = ModelShape()
LLAMA_70B = ModelShape("llama-3.1-70b", 80, 8192, 64, 8, 128, 28672, 128256)
OPT_125M_LIKE = ModelShape("opt-125m-shape", 12, 768, 12, 12, 64, 3072, 50272, 10000.0)


class DecodeRunner:
    """Owns per-rank weights + paged KV caches and runs one decode step (optionally as a CUDA graph)."""

    def __init__(self, shape: ModelShape, batch: int, max_ctx: int, *, block_size: int = 16,
                 kv_cache_dtype: str = "fp8", weight_quant: str = "fp8", tp_size: int = 1, tp_rank: int = 0,
                 device: torch.device | str = "cuda:0", seed: int = 0, num_layers: int | None = None,
                 tp_group=None, fused: bool = True, skip_collectives: bool = False):
        self.s = shape
        self.B, self.max_ctx, self.bs = batch, max_ctx, block_size
        self.kv_dtype, self.wq = kv_cache_dtype, weight_quant
        self.tp, self.rank, self.group = tp_size, tp_rank, tp_group
        # development aid: run ONE rank's shard of a TP=N model on a single GPU with every collective skipped (the sums
        # are wrong, the per-rank kernels and their cost are exactly those of the N-GPU step minus the all-reduces)
        self.skip_collectives = skip_collectives
        self.fused = fused  # use the fused norm+quant / silu+quant / rope+cache kernels
        self.dev = torch.device(device)
        self.L = num_layers if num_layers is not None else shape.num_layers
        assert shape.num_q_heads % tp_size == 0 and shape.num_kv_heads % tp_size == 0
        assert shape.intermediate % tp_size == 0 and weight_quant in ("fp8", "nvfp4")
        self.Hq, self.Hkv, self.D = shape.num_q_heads // tp_size, shape.num_kv_heads // tp_size, shape.head_size
        self.inter = shape.intermediate // tp_size
        self.vocab_shard = (shape.vocab + tp_size - 1) // tp_size
        self.dtype = torch.bfloat16
        # ONE seeded full model, identical on every rank, sharded the way vLLM's parallel linears shard a
        # checkpoint (tp_shard.py): TP=N computes the same function as TP=1 up to the quantisation / summation
        # order noise of the row-parallel all-reduce (tests/test_gpu_tp_equivalence.py, bench.py "tp_check").
        g = torch.Generator(device=self.dev)
        g.manual_seed(seed)
        H = shape.hidden
        lay = tp_shard.TPLayout(shape.num_q_heads, shape.num_kv_heads, shape.head_size, shape.intermediate, tp_size)

        def w(n, k):
            return (torch.randn(n, k, device=self.dev, generator=g, dtype=torch.float32) * 0.02).to(self.dtype)

        self.layers = []
        for _ in range(self.L):
            full = {
                "qkv": w((shape.num_q_heads + 2 * shape.num_kv_heads) * self.D, H),
                "o": w(H, shape.num_q_heads * self.D),
                "gate_up": w(2 * shape.intermediate, H),
                "down": w(H, shape.intermediate),
            }
            shard = {
                "qkv": lambda t: tp_shard.shard_qkv(t, lay, tp_rank),
                "o": lambda t: tp_shard.shard_row_parallel(t, tp_size, tp_rank),
                "gate_up": lambda t: tp_shard.shard_gate_up(t, lay, tp_rank),
                "down": lambda t: tp_shard.shard_row_parallel(t, tp_size, tp_rank),
            }
            lw = {name: self._quant_weight(full[name], shard[name]) for name in full}
            lw["ln1"] = torch.ones(H, device=self.dev, dtype=self.dtype)
            lw["ln2"] = torch.ones(H, device=self.dev, dtype=self.dtype)
            self.layers.append(lw)
            del full
        self.final_ln = torch.ones(H, device=self.dev, dtype=self.dtype)
        self.embed = w(shape.vocab, H)
        self.lm_head = tp_shard.shard_vocab_rows(w(shape.vocab, H), tp_size, tp_rank).contiguous()
        self.vocab_rows = self.lm_head.shape[0]  # the last vocab shard may be short
        # rotary cache [max_pos, rot_dim] = cos | sin
        inv = 1.0 / (shape.rope_theta ** (torch.arange(0, self.D, 2, device=self.dev).float() / self.D))
        fr = torch.outer(torch.arange(max_ctx + 8, device=self.dev).float(), inv)
        self.cos_sin = torch.cat([fr.cos(), fr.sin()], -1).to(self.dtype)
        # paged KV cache, vLLM 0.22 layout (2, num_blocks, block_size, Hkv, D) per layer
        self.blocks_per_seq = (max_ctx + block_size - 1) // block_size
        self.num_blocks = batch * self.blocks_per_seq + 1
        cdt = torch.uint8 if kv_cache_dtype.startswith("fp8") else self.dtype
        self.kv_caches = []
        for _ in range(self.L):
            # the full cache (all kv heads) from the shared generator; this rank keeps its heads (tp_shard.shard_kv_cache)
            shp = (2, self.num_blocks, block_size, shape.num_kv_heads, self.D)
            if cdt == torch.uint8:  # random finite e4m3 bytes (exponent field != 0xF)
                kv = torch.randint(0, 256, shp, device=self.dev, generator=g, dtype=torch.uint8) & 0x77
            else:
                kv = ((torch.rand(shp, device=self.dev, generator=g) * 2 - 1) * self.D ** -0.5).to(cdt)
            self.kv_caches.append(tp_shard.shard_kv_cache(kv, lay, tp_rank) if tp_size > 1 else kv)
            del kv
        self.k_scale = torch.tensor([0.05], device=self.dev)
        self.v_scale = torch.tensor([0.05], device=self.dev)
        # static step buffers (CUDA-graph inputs)
        self.input_ids = torch.zeros(batch, dtype=torch.int64, device=self.dev)
        self.positions = torch.zeros(batch, dtype=torch.int64, device=self.dev)
        self.slot_mapping = torch.zeros(batch, dtype=torch.int64, device=self.dev)
        self.seq_lens = torch.zeros(batch, dtype=torch.int32, device=self.dev)
        self.block_table = torch.zeros(batch, self.blocks_per_seq, dtype=torch.int32, device=self.dev)
        self.query_start_loc = torch.arange(batch + 1, dtype=torch.int32, device=self.dev)
        self.next_tokens = torch.zeros(batch, dtype=torch.int64, device=self.dev)
        self.num_splits = ops.paged_decode_num_splits(batch, self.Hq, self.Hkv, max_ctx, self.dev, self.D,
                                                      kv_cache_dtype if kv_cache_dtype.startswith("fp8") else "auto")
        ws = ops.paged_decode_workspace_bytes(batch, self.Hq, self.D, self.num_splits)
        self.attn_ws = torch.empty(max(ws, 16), dtype=torch.uint8, device=self.dev)
        self.scale = self.D ** -0.5
        self.graph: torch.cuda.CUDAGraph | None = None
        self.launches_per_step = 0
        # Row-parallel partial sums: two-shot all-reduce over NVLink peer memory (b200_peer_all_reduce) on
        # symmetric buffers the o_proj / down_proj GEMMs write into; NCCL when that cannot be set up.
        self.par = None
        if tp_size > 1 and not skip_collectives and os.environ.get("B200_PEER_ALLREDUCE", "1") != "0":
            try:
                from .peer_comm import PeerAllReduce

                self.par = PeerAllReduce(tp_group if tp_group is not None else dist.group.WORLD, self.dev, 2,
                                         batch * shape.hidden, self.dtype)
            except Exception as e:  # noqa: BLE001 - any failure here just means NCCL is used
                print(f"[decode_runner] peer all-reduce unavailable ({e!r}); using NCCL", file=sys.stderr, flush=True)
                self.par = None

    # ----------------------------------------------------------------- weights / linears
    def _quant_weight(self, w: torch.Tensor, shard):
        """Quantise the FULL weight (one per-tensor scale / NVFP4 global scale for the whole checkpoint tensor,
        like a pre-quantised checkpoint) and keep this rank's shard of the quantised tensor."""
        if self.wq == "fp8":
            q, s = ops.scaled_fp8_quant(w)  # per tensor, [N,K] K-contiguous == column-major [K,N]
            return {"w": shard(q.view(torch.uint8)).contiguous().view(ops.FP8_DTYPE), "s": s}
        gs = (448.0 * 6.0 / w.abs().max().float()).reshape(1)
        q, sf = ops.scaled_fp4_quant(shard(w).contiguous(), gs)  # 16-element blocks never straddle a shard boundary
        return {"w": q, "sf": sf, "gs": gs}

    def _linear(self, x: torch.Tensor, lw, out: torch.Tensor | None = None) -> torch.Tensor:
        if self.wq == "fp8":
            xq, xs = ops.scaled_fp8_quant(x, use_per_token_if_dynamic=True)
            return ops.cutlass_scaled_mm(xq, lw["w"].t(), xs, lw["s"], self.dtype, out=out)
        # NVFP4: static activation global scale (ModelOpt input_scale), here a fixed calibration constant
        self._fp4_consts(lw)
        xq, xsf = ops.scaled_fp4_quant(x, lw["ags"])
        return ops.cutlass_scaled_fp4_mm(xq, lw["w"], xsf, lw["sf"], lw["alpha"], self.dtype, out=out)

    def _fp4_consts(self, lw) -> None:
        if "ags" not in lw:
            lw["ags"] = torch.tensor([448.0 * 6.0 / 8.0], device=self.dev)
            lw["alpha"] = (1.0 / (lw["ags"] * lw["gs"])).float()

    def _fp4_bufs(self, m: int, k: int) -> tuple[torch.Tensor, torch.Tensor]:
        """Packed codes + swizzled scale buffer; torch.empty: our quantisers write every padded scale themselves."""
        return (torch.empty(m, k // 2, dtype=torch.uint8, device=self.dev),
                torch.empty(ops.round_up(m, 128), ops.round_up(k // 16, 4) // 4, dtype=torch.int32, device=self.dev))

    def _mm_fp4(self, xq: torch.Tensor, xsf: torch.Tensor, lw, out: torch.Tensor | None = None) -> torch.Tensor:
        return ops.cutlass_scaled_fp4_mm(xq, lw["w"], xsf.view(ops.FP8_DTYPE), lw["sf"], lw["alpha"], self.dtype, out=out)

    def _reduce_fp4(self, xq: torch.Tensor, xsf: torch.Tensor, lw, which: int) -> torch.Tensor:
        """Row-parallel NVFP4 linear + sum over the ranks."""
        if self.par is None:
            return self._all_reduce(self._mm_fp4(xq, xsf, lw))
        out = self.par.view(which, self.B, self.s.hidden)
        self._mm_fp4(xq, xsf, lw, out=out)
        if self.par.lamport or (self.tp <= 4 and not self.par.multicast_ptr):
            return self.par.all_reduce_out(which, torch.empty_like(out))
        self.par.all_reduce_(which)  # two-shot, in place (through the switch when the fabric has multicast)
        return out

    def _row_parallel(self, x: torch.Tensor, lw, which: int) -> torch.Tensor:
        """Row-parallel linear (quant + GEMM, either weight format) + sum over the ranks; the GEMM writes its
        partial sums straight into the symmetric buffer `which` (o_proj: 0, down_proj: 1)."""
        if self.par is None:
            return self._all_reduce(self._linear(x, lw))
        out = self.par.view(which, self.B, self.s.hidden)
        self._linear(x, lw, out=out)
        if self.tp <= 4 or self.par.lamport:
            return self.par.all_reduce_out(which, torch.empty_like(out))
        self.par.all_reduce_(which)
        return out

    def _all_reduce(self, x: torch.Tensor) -> torch.Tensor:
        if self.tp > 1 and not self.skip_collectives:
            dist.all_reduce(x, group=self.group)
        return x

    # ----------------------------------------------------------------------- the step
    def set_batch(self, seq_len: int, block_table: torch.Tensor | None = None, seed: int = 0) -> None:
        """Synthetic decode state: every sequence has `seq_len` tokens, the newest at position seq_len-1."""
        g = torch.Generator(device=self.dev)
        g.manual_seed(1234 + seed)  # same on every rank: the scheduler state is global
        if block_table is None:
            perm = torch.randperm(self.num_blocks - 1, device=self.dev, generator=g)[: self.B * self.blocks_per_seq] + 1
            block_table = perm.reshape(self.B, self.blocks_per_seq).to(torch.int32)
        self.block_table.copy_(block_table)
        self.seq_lens.fill_(seq_len)
        self.positions.fill_(seq_len - 1)
        self.input_ids.copy_(torch.randint(0, self.s.vocab, (self.B,), device=self.dev, generator=g))
        ops.compute_slot_mapping(self.query_start_loc, self.positions, self.block_table, self.bs, self.slot_mapping)

    def _mm_fp8(self, xq: torch.Tensor, xs: torch.Tensor, lw, out: torch.Tensor | None = None) -> torch.Tensor:
        return ops.cutlass_scaled_mm(xq, lw["w"].t(), xs, lw["s"], self.dtype, out=out)

    def _row_parallel_fp8(self, xq: torch.Tensor, xs: torch.Tensor, lw, which: int) -> torch.Tensor:
        """Row-parallel linear + sum over the ranks (o_proj: which = 0, down_proj: which = 1)."""
        if self.par is None:
            return self._all_reduce(self._mm_fp8(xq, xs, lw))
        out = self.par.view(which, self.B, self.s.hidden)
        self._mm_fp8(xq, xs, lw, out=out)
        # one-shot (start barrier only; o/down alternate buffers): NVLS in-switch sum at any TP when the fabric has
        # multicast, else (tp-1) MB of peer reads, which beat a second cross-GPU barrier up to TP 4
        if self.tp <= 4 or self.par.lamport:
            return self.par.all_reduce_out(which, torch.empty_like(out))
        self.par.all_reduce_(which)
        return out

    def step_eager(self) -> torch.Tensor:
        s, B = self.s, self.B
        hidden = self.embed[self.input_ids]
        residual = torch.zeros_like(hidden)
        attn_out = torch.empty(B, self.Hq, self.D, device=self.dev, dtype=self.dtype)
        fused = self.wq == "fp8" and self.fused
        # TP <= 4: the all-reduce after a row-parallel GEMM is fused with the residual add + RMSNorm + FP8 quant
        # that follows it (b200_peer_all_reduce_rms_norm_quant); `pending` = the down_proj partial sums of the
        # previous layer still sit un-reduced in peer buffer 1
        # (measured, profiles/r02_allreduce.md: at 8 ranks the fused forms lose to the in-place two-shot reduce
        # followed by the separate norm + quant kernel - 4.08 vs 3.79 ms/step - unless the barrier-free Lamport
        # protocol is on, which fuses at any TP)
        fuse_ar = fused and self.par is not None and (self.tp <= 4 or self.par.lamport)
        pending = False
        fused4 = self.wq == "nvfp4" and self.fused
        for lw, kv in zip(self.layers, self.kv_caches):
            kc, vc = kv.unbind(0)
            if fused4:
                # NVFP4, 10-11 launches per layer: norm+quant, GEMM, rope+cache, attention (+combine), quant, GEMM
                # (+ all-reduce), norm+quant, GEMM, silu*mul+quant, GEMM (+ all-reduce)
                for name in ("qkv", "o", "gate_up", "down"):
                    self._fp4_consts(lw[name])
                fuse4_ar = self.par is not None and self.par.lamport
                xq, xsf = self._fp4_bufs(B, s.hidden)
                if pending:  # the previous layer's down_proj partial sums are still un-reduced in peer buffer 1
                    self.par.all_reduce_rms_norm_nvfp4_quant(1, B, lw["ln1"], s.rms_eps, residual, lw["qkv"]["ags"],
                                                             result=xq, result_block_scale=xsf)
                else:
                    ops.rms_norm_nvfp4_quant(hidden, lw["ln1"], s.rms_eps, lw["qkv"]["ags"], residual=residual,
                                             result=xq, result_block_scale=xsf)
                qkv = self._mm_fp4(xq, xsf, lw["qkv"])
                q, k, v = qkv.split([self.Hq * self.D, self.Hkv * self.D, self.Hkv * self.D], dim=-1)
                ops.rotary_embedding_and_cache_write(self.positions, q, k, v, self.D, self.cos_sin, kc, vc,
                                                     self.slot_mapping, self.kv_dtype, self.k_scale, self.v_scale)
                ops.paged_attention_decode(attn_out, q.view(B, self.Hq, self.D), kc, vc, self.block_table,
                                           self.seq_lens, self.scale, self.max_ctx, self.kv_dtype, self.k_scale,
                                           self.v_scale, num_splits=self.num_splits, workspace=self.attn_ws)
                oq, osf = self._fp4_bufs(B, self.Hq * self.D)
                ops.scaled_fp4_quant_out(attn_out.view(B, -1), lw["o"]["ags"], True, oq, osf)
                xq, xsf = self._fp4_bufs(B, s.hidden)
                if fuse4_ar:  # o_proj partial sums -> peer buffer 0 -> all-reduce + add + RMSNorm + NVFP4 quant
                    self._mm_fp4(oq, osf, lw["o"], out=self.par.view(0, B, s.hidden))
                    self.par.all_reduce_rms_norm_nvfp4_quant(0, B, lw["ln2"], s.rms_eps, residual, lw["gate_up"]["ags"],
                                                             result=xq, result_block_scale=xsf)
                else:
                    hidden = self._reduce_fp4(oq, osf, lw["o"], 0)
                    ops.rms_norm_nvfp4_quant(hidden, lw["ln2"], s.rms_eps, lw["gate_up"]["ags"], residual=residual,
                                             result=xq, result_block_scale=xsf)
                gu = self._mm_fp4(xq, xsf, lw["gate_up"])
                aq, asf = self._fp4_bufs(B, self.inter)
                ops.silu_and_mul_nvfp4_quant(aq, asf, gu, lw["down"]["ags"])
                if fuse4_ar:
                    self._mm_fp4(aq, asf, lw["down"], out=self.par.view(1, B, s.hidden))
                    pending = True
                else:
                    hidden = self._reduce_fp4(aq, asf, lw["down"], 1)
                continue
            if fused:
                # 10 launches per layer: norm+quant, GEMM, rope+cache, attention (+combine), quant, GEMM,
                # norm+quant, GEMM, silu*mul+quant, GEMM
                if pending:
                    xq, xs = self.par.all_reduce_rms_norm_quant(1, B, lw["ln1"], s.rms_eps, residual)
                else:
                    xq, xs = ops.rms_norm_dynamic_per_token_quant(hidden, lw["ln1"], s.rms_eps, residual=residual)
                qkv = self._mm_fp8(xq, xs, lw["qkv"])
                q, k, v = qkv.split([self.Hq * self.D, self.Hkv * self.D, self.Hkv * self.D], dim=-1)
                ops.rotary_embedding_and_cache_write(self.positions, q, k, v, self.D, self.cos_sin, kc, vc,
                                                     self.slot_mapping, self.kv_dtype, self.k_scale, self.v_scale)
            else:
                ops.fused_add_rms_norm(hidden, residual, lw["ln1"], s.rms_eps)
                qkv = self._linear(hidden, lw["qkv"])
                q, k, v = qkv.split([self.Hq * self.D, self.Hkv * self.D, self.Hkv * self.D], dim=-1)
                ops.rotary_embedding(self.positions, q, k, self.D, self.cos_sin)
                ops.reshape_and_cache_flash(k.view(B, self.Hkv, self.D), v.view(B, self.Hkv, self.D), kc, vc,
                                            self.slot_mapping, self.kv_dtype, self.k_scale, self.v_scale)
            ops.paged_attention_decode(attn_out, q.view(B, self.Hq, self.D), kc, vc, self.block_table, self.seq_lens,
                                       self.scale, self.max_ctx, self.kv_dtype, self.k_scale, self.v_scale,
                                       num_splits=self.num_splits, workspace=self.attn_ws)
            if fuse_ar:
                oq, o_s = ops.scaled_fp8_quant(attn_out.view(B, -1), use_per_token_if_dynamic=True)
                self._mm_fp8(oq, o_s, lw["o"], out=self.par.view(0, B, s.hidden))
                xq, xs = self.par.all_reduce_rms_norm_quant(0, B, lw["ln2"], s.rms_eps, residual)
                gu = self._mm_fp8(xq, xs, lw["gate_up"])
                aq, a_s = ops.silu_and_mul_dynamic_per_token_quant(gu)
                self._mm_fp8(aq, a_s, lw["down"], out=self.par.view(1, B, s.hidden))
                pending = True
                continue
            if fused and self.par is not None:
                oq, o_s = ops.scaled_fp8_quant(attn_out.view(B, -1), use_per_token_if_dynamic=True)
                hidden = self._row_parallel_fp8(oq, o_s, lw["o"], 0)
            else:
                hidden = self._row_parallel(attn_out.view(B, -1), lw["o"], 0)
            if fused:
                xq, xs = ops.rms_norm_dynamic_per_token_quant(hidden, lw["ln2"], s.rms_eps, residual=residual)
                gu = self._mm_fp8(xq, xs, lw["gate_up"])
                aq, a_s = ops.silu_and_mul_dynamic_per_token_quant(gu)
                hidden = self._row_parallel_fp8(aq, a_s, lw["down"], 1)
            else:
                ops.fused_add_rms_norm(hidden, residual, lw["ln2"], s.rms_eps)
                gu = self._linear(hidden, lw["gate_up"])
                act = torch.empty(B, self.inter, device=self.dev, dtype=self.dtype)
                ops.silu_and_mul(act, gu)
                hidden = self._row_parallel(act, lw["down"], 1)
        if pending:  # the last layer's down_proj: plain one-shot all-reduce, then the final norm
            hidden = self.par.all_reduce_out(1, torch.empty(B, s.hidden, device=self.dev, dtype=self.dtype))
        ops.fused_add_rms_norm(hidden, residual, self.final_ln, s.rms_eps)
        self.last_hidden = hidden  # [B, hidden] after the final norm (TP-equivalence checks read it)
        logits = torch.matmul(hidden, self.lm_head.t())  # plain library GEMM (bf16 lm_head, as in vLLM)
        if self.tp == 1 or self.skip_collectives:
            self.next_tokens.copy_(logits.argmax(-1))
        else:
            mx, idx = logits.float().max(-1)
            both = torch.stack([mx, (idx + self.rank * self.vocab_shard).float()], -1)
            gathered = [torch.empty_like(both) for _ in range(self.tp)]
            dist.all_gather(gathered, both, group=self.group)
            allv = torch.stack(gathered, 0)
            win = allv[..., 0].argmax(0)
            self.next_tokens.copy_(allv[win, torch.arange(B, device=self.dev), 1].long())
        return self.next_tokens

    def capture(self, warmup: int = 2) -> None:
        """Capture the step into a CUDA graph (what vLLM's FULL cudagraph mode does for decode)."""
        side = torch.cuda.Stream(self.dev)
        side.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self.step_eager()
        torch.cuda.current_stream(self.dev).wait_stream(side)
        torch.cuda.synchronize(self.dev)
        before = _lib.num_kernels_launched()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.step_eager()
        self.launches_per_step = _lib.num_kernels_launched() - before

    def step(self) -> torch.Tensor:
        if self.graph is not None:
            self.graph.replay()
            return self.next_tokens
        return self.step_eager()

    # --------------------------------------------------------------------- accounting
    def attention_bytes_per_layer(self, seq_len: int) -> int:
        """sum_s len_s * Hkv * D * 2 * kv_itemsize + B * Hq * D * (2 + 2)."""
        it = 1 if self.kv_dtype.startswith("fp8") else 2
        return self.B * seq_len * self.Hkv * self.D * 2 * it + self.B * self.Hq * self.D * 4

    def weight_bytes_per_layer(self) -> int:
        n = sum(lw["w"].numel() * lw["w"].element_size() for lw in self.layers[0].values() if isinstance(lw, dict))
        return n


TP_CHECK_SHAPE = ModelShape("tp-check", 2, 1024, 16, 8, 128, 4096, 8192)


def tp_equivalence_check(tp_size: int, tp_rank: int, group, device, *, batch: int = 16, ctx: int = 500,
                         shape: ModelShape = TP_CHECK_SHAPE, weight_quant: str = "fp8") -> dict:
    """Does the TP=N step compute the same function as TP=1?  Every rank builds the TP=N runner (its shard of
    ONE seeded model) and, locally, the unsharded TP=1 runner of the same model, runs one eager step of each and
    compares the hidden state after the final norm and the greedy tokens.  Differences come only from where the
    dynamic per-token activation scales are taken (per shard vs whole row) and the all-reduce's summation
    order, so the bound is a few fp8 ulps of the activations, not bit equality."""
    one = DecodeRunner(shape, batch, ctx, tp_size=1, tp_rank=0, device=device, weight_quant=weight_quant)
    one.set_batch(ctx)
    t1 = one.step_eager().clone()
    h1 = one.last_hidden.float().clone()
    many = DecodeRunner(shape, batch, ctx, tp_size=tp_size, tp_rank=tp_rank, device=device, tp_group=group,
                        weight_quant=weight_quant)
    many.set_batch(ctx)
    tn = many.step_eager().clone()
    hn = many.last_hidden.float()
    torch.cuda.synchronize(device)
    rel = float((hn - h1).norm() / h1.norm())
    # a sharded-weights mistake (wrong head / column slice) gives rel ~ 1.4 (uncorrelated vectors)
    return {"tp": tp_size, "hidden_rel_l2": rel, "hidden_max_abs": float((hn - h1).abs().max()),
            "tokens_equal_frac": float((tn == t1).float().mean()), "peer_allreduce": many.par is not None,
            "weight_quant": weight_quant}
