"""Tensor-parallel sharding of the decode hot path's weights and KV heads (host logic, no kernels).

Mirrors what the reference's parallel linear layers do to a checkpoint at load time:
  QKVParallelLinear       vllm/model_executor/layers/linear.py  - q/k/v rows split by HEADS, each rank keeps
                          its q heads and its kv heads ([q_r | k_r | v_r] fused output rows)
  MergedColumnParallel    same file - gate and up each split by rows, fused as [gate_r | up_r]
  RowParallelLinear       same file - o_proj / down_proj split by INPUT columns; the partial outputs are summed
                          with an all-reduce (NCCL over NVLink in `DecodeRunner`, gloo in the CPU tests)
and how the paged KV cache is sharded: rank r owns kv heads [r*Hkv/tp, (r+1)*Hkv/tp) of every block.

The path shards with exactly one exchange step per row-parallel linear (the all-reduce after o_proj and
down_proj); attention itself needs no collective because a kv head and its query-head group live on one rank.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch


@dataclass(frozen=True)
class TPLayout:
    num_q_heads: int
    num_kv_heads: int
    head_size: int
    intermediate: int
    tp_size: int

    def __post_init__(self):
        if self.tp_size < 1:
            raise ValueError("tp_size must be >= 1")
        if self.num_q_heads % self.tp_size or self.intermediate % self.tp_size:
            raise ValueError("num_q_heads and intermediate must be divisible by tp_size")
        if self.num_kv_heads % self.tp_size:
            # the reference replicates kv heads when tp_size > num_kv_heads; this path does not build that
            raise ValueError("num_kv_heads must be divisible by tp_size (kv-head replication is not built)")

    @property
    def q_heads_per_rank(self) -> int:
        return self.num_q_heads // self.tp_size

    @property
    def kv_heads_per_rank(self) -> int:
        return self.num_kv_heads // self.tp_size

    def q_head_range(self, rank: int) -> range:
        n = self.q_heads_per_rank
        return range(rank * n, (rank + 1) * n)

    def kv_head_range(self, rank: int) -> range:
        n = self.kv_heads_per_rank
        return range(rank * n, (rank + 1) * n)


def shard_qkv(w_qkv: torch.Tensor, lay: TPLayout, rank: int) -> torch.Tensor:
    """[ (Hq + 2 Hkv) * D, hidden ] fused q|k|v rows -> this rank's [ (Hq_r + 2 Hkv_r) * D, hidden ]."""
    D, Hq, Hkv = lay.head_size, lay.num_q_heads, lay.num_kv_heads
    if w_qkv.shape[0] != (Hq + 2 * Hkv) * D:
        raise ValueError(f"qkv weight has {w_qkv.shape[0]} rows, expected {(Hq + 2 * Hkv) * D}")
    q, k, v = w_qkv.split([Hq * D, Hkv * D, Hkv * D], dim=0)
    qr, kr = lay.q_head_range(rank), lay.kv_head_range(rank)
    return torch.cat([q[qr.start * D:qr.stop * D], k[kr.start * D:kr.stop * D], v[kr.start * D:kr.stop * D]], dim=0)


def shard_gate_up(w_gate_up: torch.Tensor, lay: TPLayout, rank: int) -> torch.Tensor:
    """[2 * intermediate, hidden] fused gate|up rows -> this rank's [2 * intermediate / tp, hidden]."""
    I = lay.intermediate
    if w_gate_up.shape[0] != 2 * I:
        raise ValueError(f"gate_up weight has {w_gate_up.shape[0]} rows, expected {2 * I}")
    n = I // lay.tp_size
    gate, up = w_gate_up.split([I, I], dim=0)
    return torch.cat([gate[rank * n:(rank + 1) * n], up[rank * n:(rank + 1) * n]], dim=0)


def shard_row_parallel(w: torch.Tensor, tp_size: int, rank: int) -> torch.Tensor:
    """[out, in] -> this rank's input-column slice [out, in / tp] (o_proj: in = Hq*D, down_proj: in = intermediate)."""
    if w.shape[1] % tp_size:
        raise ValueError("input dimension must be divisible by tp_size")
    n = w.shape[1] // tp_size
    return w[:, rank * n:(rank + 1) * n].contiguous()


def shard_kv_cache(kv_cache: torch.Tensor, lay: TPLayout, rank: int) -> torch.Tensor:
    """(2, num_blocks, block_size, Hkv, D) -> the rank's kv heads (2, num_blocks, block_size, Hkv_r, D)."""
    r = lay.kv_head_range(rank)
    return kv_cache[:, :, :, r.start:r.stop, :].contiguous()


def shard_vocab_rows(w: torch.Tensor, tp_size: int, rank: int) -> torch.Tensor:
    """lm_head / embedding rows split like VocabParallelEmbedding (last shard may be short)."""
    n = (w.shape[0] + tp_size - 1) // tp_size
    return w[rank * n:min(w.shape[0], (rank + 1) * n)]
