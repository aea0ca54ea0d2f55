"""b200_peer_all_reduce / b200_peer_all_reduce_out against NCCL on two GPUs (skipped on a single-GPU box)."""
import os
import socket
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from vllm_b200.peer_comm import PeerAllReduce

    B, H = 48, 1024
    par = PeerAllReduce(dist.group.WORLD, dev, 2, B * H)
    g = torch.Generator(device=dev).manual_seed(7 + rank)
    ok = True
    for it in range(6):
        x = torch.randn(B, H, device=dev, generator=g).to(torch.bfloat16)
        ref = x.float().clone()
        dist.all_reduce(ref)  # fp32 sum of the bf16 inputs
        par.view(it & 1, B, H).copy_(x)
        if it % 3 == 2:
            par.all_reduce_(it & 1)
            got = par.view(it & 1, B, H)
        else:
            got = par.all_reduce_out(it & 1, torch.empty(B, H, device=dev, dtype=torch.bfloat16))
        torch.cuda.synchronize()
        ok = ok and torch.equal(got, ref.to(torch.bfloat16))  # fp32 accumulation, one final rounding: exact
        allg = [torch.empty_like(got) for _ in range(world)]
        dist.all_gather(allg, got.contiguous())
        ok = ok and all(torch.equal(allg[0], t) for t in allg)  # identical on every rank
    # fused all-reduce + residual add + RMSNorm + per-token FP8 quant == the two separate kernels, bit for bit
    from vllm_b200 import ops

    w = (torch.rand(H, device=dev, generator=g) + 0.5).to(torch.bfloat16)
    dist.broadcast(w, 0)
    for it in range(4):
        x = torch.randn(B, H, device=dev, generator=g).to(torch.bfloat16)
        res = torch.randn(B, H, device=dev, generator=torch.Generator(device=dev).manual_seed(100 + it)).to(torch.bfloat16)
        par.view(it & 1, B, H).copy_(x)
        summed = par.all_reduce_out(it & 1, torch.empty(B, H, device=dev, dtype=torch.bfloat16)).clone()
        res_ref = res.clone()
        xq_ref, xs_ref = ops.rms_norm_dynamic_per_token_quant(summed, w, 1e-5, residual=res_ref)
        it2 = (it + 1) & 1
        par.view(it2, B, H).copy_(x)
        res_f = res.clone()
        xq, xs = par.all_reduce_rms_norm_quant(it2, B, w, 1e-5, res_f)
        torch.cuda.synchronize()
        ok = ok and torch.equal(xq.view(torch.uint8), xq_ref.view(torch.uint8)) and torch.equal(xs, xs_ref)
        ok = ok and torch.equal(res_f, res_ref)
    if rank == 0:
        q.put(ok)
    torch.cuda.synchronize()
    dist.barrier()
    os._exit(0)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_all_reduce_matches_fp32_sum():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    assert q.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
