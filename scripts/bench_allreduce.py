#!/usr/bin/env python
"""All-reduce of the decode step's row-parallel output: NCCL vs the b200 peer-memory / NVLS kernels, each captured
32x into a CUDA graph and replayed; max over ranks.  Sizes: 128 x 4096 bf16 = 1 MB (configs[2]) and 256 x 8192 bf16 =
4 MB (configs[3]).  "lamport": all_reduce_out / the fused form through the barrier-free protocol (for it "one_shot_us"
is the out-of-place all-reduce, one-shot up to 4 ranks and two-shot above).  Run under torchrun:
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_allreduce.py [--json F]"""
import json, os, sys
from pathlib import Path
import torch
import torch.distributed as dist
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
os.environ.setdefault("B200_NVLS", "force")  # map the multicast pointer whenever the fabric has one; toggled below
from vllm_b200 import ops
from vllm_b200.peer_comm import PeerAllReduce

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)


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
    del g
    return round(float(t.item()), 2)


rows = []
for B, H in ((128, 4096), (256, 8192)):
    par = PeerAllReduce(dist.group.WORLD, dev, 2, B * H)
    mc_real = par.multicast_ptr
    x = torch.randn(B, H, device=dev).to(torch.bfloat16)
    ref = x.float().clone(); dist.all_reduce(ref)          # fp32 sum of the bf16 inputs
    w = (torch.rand(H, device=dev) + 0.5).to(torch.bfloat16); dist.broadcast(w, 0)
    res0 = torch.randn(B, H, device=dev).to(torch.bfloat16); dist.broadcast(res0, 0)
    r = {"world": world, "bytes": B * H * 2, "nvls_available": bool(mc_real)}
    y = x.clone()
    r["nccl_us"] = timed(lambda: dist.all_reduce(y))
    o1 = torch.empty(B, H, device=dev, dtype=torch.bfloat16)
    flip = [0]

    def one():
        par.all_reduce_out(flip[0], o1); flip[0] ^= 1

    resid = res0.clone()

    def fused():
        par.all_reduce_rms_norm_quant(flip[0], B, w, 1e-5, resid); flip[0] ^= 1

    has_lp = par.lamport
    for name, mc, lp in (("peer", 0, False), ("nvls", mc_real, False), ("lamport", 0, True)):
        if (name == "nvls" and not mc_real) or (lp and not has_lp):
            continue
        par.multicast_ptr = mc
        par.lamport = lp
        # correctness of each form first
        par.view(0, B, H).copy_(x); par.view(1, B, H).copy_(x); torch.cuda.synchronize(); dist.barrier()
        par.all_reduce_(0); torch.cuda.synchronize()
        e2 = (par.view(0, B, H).float() - ref).abs().max().item()
        flip[0] = 1; one(); torch.cuda.synchronize()
        e1 = (o1.float() - ref).abs().max().item()
        # fused form == all_reduce_out + the stand-alone fused norm/quant kernel of the same flavour, bit for bit
        par.view(0, B, H).copy_(x); par.view(1, B, H).copy_(x); torch.cuda.synchronize(); dist.barrier()
        summed = par.all_reduce_out(0, torch.empty_like(o1)).clone()
        rr = res0.clone(); xq_ref, xs_ref = ops.rms_norm_dynamic_per_token_quant(summed, w, 1e-5, residual=rr)
        rf = res0.clone(); xq, xs = par.all_reduce_rms_norm_quant(1, B, w, 1e-5, rf); torch.cuda.synchronize()
        fused_ok = bool(torch.equal(xq.view(torch.uint8), xq_ref.view(torch.uint8)) and torch.equal(xs, xs_ref)
                        and torch.equal(rf, rr))
        ok = torch.tensor([int(fused_ok)], device=dev); dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        torch.cuda.synchronize(); dist.barrier()
        r[name] = {"two_shot_us": None if lp else timed(lambda: par.all_reduce_(0)), "one_shot_us": timed(one),
                   "fused_norm_quant_us": timed(fused), "err_two_shot": e2, "err_one_shot": e1,
                   "fused_bit_identical_all_ranks": bool(ok.item())}
    par.multicast_ptr = mc_real
    par.lamport = has_lp
    # what the fused kernel replaces when the reduce is not fused: the stand-alone norm + quant kernel
    hid = torch.randn(B, H, device=dev).to(torch.bfloat16)
    r["rms_norm_quant_alone_us"] = timed(lambda: ops.rms_norm_dynamic_per_token_quant(hid, w, 1e-5, residual=resid))
    rows.append(r)
    if rank == 0:
        print(json.dumps(r), flush=True)
    del par
if rank == 0 and "--json" in sys.argv:
    Path(sys.argv[sys.argv.index("--json") + 1]).write_text("\n".join(json.dumps(r) for r in rows) + "\n")
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
os._exit(0)
