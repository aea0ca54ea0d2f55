#!/usr/bin/env python
"""All-reduce of the decode step's row-parallel output (128 x 4096 bf16 = 1 MB): NCCL vs b200_peer_all_reduce,
each captured 32x into a CUDA graph and replayed.  Run under torchrun:
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_allreduce.py"""
import os, sys
from pathlib import Path
import torch
import torch.distributed as dist
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vllm_b200.peer_comm import PeerAllReduce

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
B, H = 128, 4096
par = PeerAllReduce(dist.group.WORLD, dev, 2, B * H)
x = torch.randn(B, H, device=dev).to(torch.bfloat16)
# correctness: same input through both
ref = x.clone(); dist.all_reduce(ref)
v = par.view(0, B, H); v.copy_(x); par.all_reduce_(0); torch.cuda.synchronize()
err = (v.float() - ref.float()).abs().max().item()
again = par.view(1, B, H); again.copy_(x); par.all_reduce_(1); torch.cuda.synchronize()
same = torch.equal(again, v)

def timed(fn, reps=32, iters=20):
    s = torch.cuda.Stream(dev)
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.synchronize(); dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / (iters * reps)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), g

y = x.clone()
nccl_us, g1 = timed(lambda: dist.all_reduce(y))
peer_us, g2 = timed(lambda: par.all_reduce_(0))
o1 = torch.empty(B, H, device=dev, dtype=torch.bfloat16)
par.view(0, B, H).copy_(x); par.view(1, B, H).copy_(x); torch.cuda.synchronize(); dist.barrier()
flip = [0]
def one():
    par.all_reduce_out(flip[0], o1); flip[0] ^= 1
one(); torch.cuda.synchronize()
err1 = (o1.float() - ref.float()).abs().max().item()
one_us, g3 = timed(one)   # NVLS (multimem.ld_reduce) when the fabric has multicast, else peer loads
nvls = bool(par.multicast_ptr)
peer_one_us = None
if nvls:
    mc, par.multicast_ptr = par.multicast_ptr, 0
    peer_one_us, g4 = timed(one)
    par.multicast_ptr = mc
if rank == 0:
    extra = f"one-shot peer loads {peer_one_us:.1f} us  " if peer_one_us is not None else ""
    print(f"world {world}: 1 MB bf16 all-reduce  NCCL {nccl_us:.1f} us  peer two-shot {peer_us:.1f} us  "
          f"one-shot{' NVLS' if nvls else ''} {one_us:.1f} us (err {err1:.4f})  {extra}"
          f"max|peer - nccl| {err:.4f}  run-to-run identical {same}  (signal pad {par._pad_bytes} B)", flush=True)
del g1, g2, g3
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
os._exit(0)
