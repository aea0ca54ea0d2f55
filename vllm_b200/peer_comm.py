"""Peer-memory all-reduce for tensor-parallel decode (host side).

`PeerAllReduce` owns one symmetric-memory allocation per rank (torch.distributed._symmetric_memory: every rank can
address every other rank's buffer over NVLink) and runs `b200_peer_all_reduce` on views of it.  The row-parallel
GEMMs write their partial sums straight into the views, the kernel reduces in place, and the consumer reads the
same view - no staging copies.  Mirrors what the reference's `CustomAllreduce`
(vllm/distributed/device_communicators/custom_all_reduce.py) does for small messages.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.distributed as dist

from . import _lib


class PeerAllReduce:
    def __init__(self, group, device: torch.device, num_buffers: int, numel: int, dtype: torch.dtype = torch.bfloat16):
        import torch.distributed._symmetric_memory as symm_mem

        if dtype not in (torch.bfloat16, torch.float16):
            raise RuntimeError("PeerAllReduce: 16-bit floating point tensors only")
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if numel % (8 * self.world):
            raise RuntimeError(f"PeerAllReduce: numel {numel} must be a multiple of 8 * world_size")
        self.dtype, self.numel, self.num_buffers = dtype, numel, num_buffers
        # The barrier flags live in the SAME symmetric allocation, behind the buffers: 128 bytes per CTA, room for
        # 256 CTAs (torch's own signal pad only has room for 72, which capped the fused kernel at 64 rows in flight).
        self._pad_bytes = 256 * 128
        self.storage = symm_mem.empty(num_buffers * numel + self._pad_bytes // 2, dtype=dtype, device=device)
        self.handle = symm_mem.rendezvous(self.storage, group)
        self.storage.zero_()
        bufs = [int(p) for p in self.handle.buffer_ptrs]
        pads = [p + num_buffers * numel * 2 for p in bufs]
        self._bufs = (C.c_uint64 * self.world)(*bufs)
        self._pads = (C.c_uint64 * self.world)(*pads)
        # NVLS: the multicast mapping of the same allocation (0 = no multicast support on this fabric / disabled)
        # B200_NVLS: "auto" (default) = above 4 ranks only - measured on B200 (profiles/r02_allreduce.md): at 2 ranks
        # one multimem.ld_reduce pass costs 13.6 us against 12.2 us for plain peer loads, at 8 ranks it replaces 7 MB
        # of peer reads per rank; "force" = whenever the fabric has multicast; "0" = never.
        mc = 0
        mode = os.environ.get("B200_NVLS", "auto")
        if mode not in ("0", "off") and (mode in ("force", "1") or self.world > 4):
            try:
                mc = int(self.handle.multicast_ptr or 0)
            except Exception:  # noqa: BLE001 - torch build / fabric without multicast support
                mc = 0
        self.multicast_ptr = mc
        self._dt = 2 if dtype == torch.bfloat16 else 1  # B200_DT_BF16 / B200_DT_F16
        # Barrier-free (Lamport) protocol: its own symmetric region (state + 3 generations of receive buffers,
        # pre-filled with the sentinel).  B200_AR_LAMPORT: "1" (default) = used by all_reduce_out and
        # all_reduce_rms_norm_quant, "0" = the flag-barrier kernels above.
        self.lamport = False
        if os.environ.get("B200_AR_LAMPORT", "1") not in ("0", "off"):
            lib = _lib.load()
            self._lp_bytes = int(lib.b200_lamport_workspace_bytes(self.world, numel))
            self._lp_storage = symm_mem.empty((self._lp_bytes + 1) // 2, dtype=dtype, device=device)
            self._lp_handle = symm_mem.rendezvous(self._lp_storage, group)
            from .ops import _stream

            _lib.check(lib.b200_lamport_workspace_init(self._lp_storage.data_ptr(), self._lp_bytes,
                                                       _stream(self._lp_storage)), "lamport_workspace_init")
            self._lp_ws = (C.c_uint64 * self.world)(*[int(p) for p in self._lp_handle.buffer_ptrs])
            self.lamport = True
        torch.cuda.synchronize(device)
        dist.barrier(group)

    def view(self, index: int, *shape: int) -> torch.Tensor:
        """The `index`-th buffer as a tensor of `shape` (the GEMM's `out=`)."""
        n = 1
        for d in shape:
            n *= d
        if n != self.numel or not 0 <= index < self.num_buffers:
            raise RuntimeError("PeerAllReduce.view: shape/index does not match the allocation")
        return self.storage[index * self.numel:(index + 1) * self.numel].view(*shape)

    def all_reduce_out(self, index: int, out: torch.Tensor) -> torch.Tensor:
        """One-shot sum of buffer `index` over the ranks into the local tensor `out`.  Consecutive calls must
        alternate between buffers (see b200_peer_all_reduce_out)."""
        from .ops import _stream

        if out.numel() != self.numel or out.dtype != self.dtype or not out.is_contiguous():
            raise RuntimeError("PeerAllReduce.all_reduce_out: out must be a contiguous tensor of the buffer's size/dtype")
        if self.lamport:
            cols = out.shape[-1] if out.dim() >= 2 and out.shape[-1] % (8 * self.world) == 0 else 8 * self.world
            src = self.storage[index * self.numel:(index + 1) * self.numel]
            rc = _lib.load().b200_lamport_all_reduce(out.data_ptr(), src.data_ptr(), self._lp_ws, self._lp_bytes, self.rank,
                                                     self.world, self.numel // cols, cols, self._dt, _stream(self.storage))
            _lib.check(rc, "lamport_all_reduce")
            return out
        if self.multicast_ptr:
            rc = _lib.load().b200_nvls_all_reduce_out(out.data_ptr(), self.multicast_ptr, self._bufs, self._pads, self.rank,
                                                      self.world, self.numel, self._dt, index * self.numel * 2,
                                                      self._pad_bytes, _stream(self.storage))
            _lib.check(rc, "nvls_all_reduce_out")
            return out
        rc = _lib.load().b200_peer_all_reduce_out(out.data_ptr(), self._bufs, self._pads, self.rank, self.world,
                                                  self.numel, self._dt, index * self.numel * 2, self._pad_bytes,
                                                  _stream(self.storage))
        _lib.check(rc, "peer_all_reduce_out")
        return out

    def all_reduce_rms_norm_quant(self, index: int, num_tokens: int, weight: torch.Tensor, epsilon: float,
                                  residual: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
        """residual += sum over ranks of buffer `index` ([num_tokens, hidden]); returns the per-token FP8 quant of
        RMSNorm(residual) * weight and its scales - one kernel (b200_peer_all_reduce_rms_norm_quant)."""
        from .ops import _stream

        hidden = self.numel // num_tokens
        if hidden * num_tokens != self.numel or residual.shape != (num_tokens, hidden) or not residual.is_contiguous():
            raise RuntimeError("PeerAllReduce.all_reduce_rms_norm_quant: shapes do not match the buffer")
        xq = torch.empty(num_tokens, hidden, dtype=torch.float8_e4m3fn, device=residual.device)
        xs = torch.empty(num_tokens, 1, dtype=torch.float32, device=residual.device)
        if self.lamport:
            src = self.storage[index * self.numel:(index + 1) * self.numel]
            rc = _lib.load().b200_lamport_all_reduce_rms_norm_quant(
                xq.data_ptr(), xs.data_ptr(), src.data_ptr(), self._lp_ws, self._lp_bytes, self.rank, self.world,
                weight.data_ptr(), residual.data_ptr(), float(epsilon), num_tokens, hidden, self._dt, _stream(self.storage))
            _lib.check(rc, "lamport_all_reduce_rms_norm_quant")
            return xq, xs
        if self.multicast_ptr:
            rc = _lib.load().b200_nvls_all_reduce_rms_norm_quant(
                xq.data_ptr(), xs.data_ptr(), self.multicast_ptr, self._bufs, self._pads, self.rank, self.world,
                index * self.numel * 2, self._pad_bytes, weight.data_ptr(), residual.data_ptr(), float(epsilon),
                num_tokens, hidden, self._dt, _stream(self.storage))
            _lib.check(rc, "nvls_all_reduce_rms_norm_quant")
            return xq, xs
        rc = _lib.load().b200_peer_all_reduce_rms_norm_quant(
            xq.data_ptr(), xs.data_ptr(), self._bufs, self._pads, self.rank, self.world, index * self.numel * 2,
            self._pad_bytes, weight.data_ptr(), residual.data_ptr(), float(epsilon), num_tokens, hidden, self._dt,
            _stream(self.storage))
        _lib.check(rc, "peer_all_reduce_rms_norm_quant")
        return xq, xs

    def all_reduce_rms_norm_nvfp4_quant(self, index: int, num_tokens: int, weight: torch.Tensor, epsilon: float,
                                        residual: torch.Tensor, global_scale: torch.Tensor,
                                        result: torch.Tensor | None = None,
                                        result_block_scale: torch.Tensor | None = None):
        """residual += sum over ranks of buffer `index`; returns the NVFP4 quant (packed codes, swizzled scale buffer)
        of RMSNorm(residual) * weight - one kernel (Lamport protocol only)."""
        from .ops import _stream, round_up

        if not self.lamport:
            raise RuntimeError("PeerAllReduce.all_reduce_rms_norm_nvfp4_quant needs the Lamport protocol (B200_AR_LAMPORT)")
        hidden = self.numel // num_tokens
        if hidden * num_tokens != self.numel or residual.shape != (num_tokens, hidden) or not residual.is_contiguous():
            raise RuntimeError("PeerAllReduce.all_reduce_rms_norm_nvfp4_quant: shapes do not match the buffer")
        if result is None:
            result = torch.empty(num_tokens, hidden // 2, dtype=torch.uint8, device=residual.device)
        if result_block_scale is None:
            result_block_scale = torch.empty(round_up(num_tokens, 128), round_up(hidden // 16, 4) // 4, dtype=torch.int32,
                                             device=residual.device)
        src = self.storage[index * self.numel:(index + 1) * self.numel]
        rc = _lib.load().b200_lamport_all_reduce_rms_norm_nvfp4_quant(
            result.data_ptr(), result_block_scale.data_ptr(), src.data_ptr(), self._lp_ws, self._lp_bytes, self.rank,
            self.world, weight.data_ptr(), residual.data_ptr(), global_scale.data_ptr(), float(epsilon), num_tokens, hidden,
            self._dt, _stream(self.storage))
        _lib.check(rc, "lamport_all_reduce_rms_norm_nvfp4_quant")
        return result, result_block_scale

    def all_reduce_(self, index: int) -> None:
        """Sum buffer `index` over the ranks, in place, on the current stream."""
        from .ops import _stream

        if self.multicast_ptr:  # two-shot through the switch: multimem.ld_reduce of my slice + one multicast store
            rc = _lib.load().b200_nvls_all_reduce(self.multicast_ptr, self._bufs, self._pads, self.rank, self.world,
                                                  self.numel, self._dt, index * self.numel * 2, self._pad_bytes,
                                                  _stream(self.storage))
            _lib.check(rc, "nvls_all_reduce")
            return
        rc = _lib.load().b200_peer_all_reduce(self._bufs, self._pads, self.rank, self.world, self.numel, self._dt,
                                              index * self.numel * 2, self._pad_bytes, _stream(self.storage))
        _lib.check(rc, "peer_all_reduce")
