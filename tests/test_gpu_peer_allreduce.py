"""b200_peer_all_reduce / _out / _rms_norm_quant.

Two ways, both through the C ABI:
  * single-GPU emulation (rank = -1, include/b200_vllm.h): all N ranks' buffers on one device, ONE cooperative
    launch with blockIdx.y = rank.  Runs on a one-GPU box, so the two-shot kernel at 8 ranks and the fused
    all-reduce + RMSNorm + quant at 4 ranks are checked wherever the GPU tests run;
  * real ranks (one process per GPU, torch symmetric memory) for world in {2, 4, 8}; skipped when the box has
    fewer GPUs.
The expected value is the fp32 sum of the bf16 inputs rounded once (the kernels add in rank order with fp32
accumulation), so the comparison is exact.
"""
import ctypes as C
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

PAD_BYTES = 64 * 32 * 4  # 64 CTAs x 32 words


def _emu_tables(world, numel, dev):
    bufs = [torch.zeros(2 * numel, dtype=torch.bfloat16, device=dev) for _ in range(world)]
    pads = [torch.zeros(PAD_BYTES // 4, dtype=torch.int32, device=dev) for _ in range(world)]
    tb = (C.c_uint64 * world)(*[b.data_ptr() for b in bufs])
    tp = (C.c_uint64 * world)(*[p.data_ptr() for p in pads])
    return bufs, pads, tb, tp


@pytest.mark.parametrize("world", [2, 4, 8])
def test_emulated_all_reduce_two_shot_and_one_shot(cuda_lib, world):
    from vllm_b200 import _lib

    dev = torch.device("cuda", 0)
    B, H = 48, 1024
    numel = B * H
    bufs, pads, tb, tp = _emu_tables(world, numel, dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    g = torch.Generator(device=dev).manual_seed(11)
    for it in range(6):
        idx = it & 1
        xs = [torch.randn(B, H, device=dev, generator=g).to(torch.bfloat16) for _ in range(world)]
        ref = torch.stack([x.float() for x in xs]).sum(0).to(torch.bfloat16)
        for b, x in zip(bufs, xs):
            b[idx * numel:(idx + 1) * numel].copy_(x.reshape(-1))
        if it % 3 == 2:  # two-shot, in place
            rc = cuda_lib.b200_peer_all_reduce(tb, tp, -1, world, numel, 2, idx * numel * 2, PAD_BYTES, st)
            _lib.check(rc, "peer_all_reduce(emu)")
            torch.cuda.synchronize()
            for b in bufs:
                assert torch.equal(b[idx * numel:(idx + 1) * numel].view(B, H), ref)
        else:  # one-shot into per-rank outputs
            out = torch.empty(world, B, H, dtype=torch.bfloat16, device=dev)
            rc = cuda_lib.b200_peer_all_reduce_out(out.data_ptr(), tb, tp, -1, world, numel, 2, idx * numel * 2,
                                                   PAD_BYTES, st)
            _lib.check(rc, "peer_all_reduce_out(emu)")
            torch.cuda.synchronize()
            for r in range(world):
                assert torch.equal(out[r], ref)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_emulated_all_reduce_rms_norm_quant(cuda_lib, world):
    """Fused all-reduce + residual add + RMSNorm + per-token FP8 quant == the separate kernels, bit for bit."""
    from vllm_b200 import _lib, ops

    dev = torch.device("cuda", 0)
    B, H = 40, 2048
    numel = B * H
    bufs, pads, tb, tp = _emu_tables(world, numel, dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    g = torch.Generator(device=dev).manual_seed(5)
    w = (torch.rand(H, device=dev, generator=g) + 0.5).to(torch.bfloat16)
    for it in range(4):
        idx = it & 1
        xs = [torch.randn(B, H, device=dev, generator=g).to(torch.bfloat16) for _ in range(world)]
        res = torch.randn(B, H, device=dev, generator=g).to(torch.bfloat16)
        summed = torch.stack([x.float() for x in xs]).sum(0).to(torch.bfloat16)
        res_ref = res.clone()
        xq_ref, xs_ref = ops.rms_norm_dynamic_per_token_quant(summed, w, 1e-5, residual=res_ref)
        for b, x in zip(bufs, xs):
            b[idx * numel:(idx + 1) * numel].copy_(x.reshape(-1))
        xq = torch.empty(world, B, H, dtype=torch.float8_e4m3fn, device=dev)
        sc = torch.empty(world, B, dtype=torch.float32, device=dev)
        resid = res.unsqueeze(0).repeat(world, 1, 1).contiguous()
        rc = cuda_lib.b200_peer_all_reduce_rms_norm_quant(xq.data_ptr(), sc.data_ptr(), tb, tp, -1, world,
                                                          idx * numel * 2, PAD_BYTES, w.data_ptr(), resid.data_ptr(),
                                                          1e-5, B, H, 2, st)
        _lib.check(rc, "peer_all_reduce_rms_norm_quant(emu)")
        torch.cuda.synchronize()
        for r in range(world):
            assert torch.equal(xq[r].view(torch.uint8), xq_ref.view(torch.uint8))
            assert torch.equal(sc[r], xs_ref.reshape(-1))
            assert torch.equal(resid[r], res_ref)


def _lamport_ws(world, max_numel, dev, lib):
    from vllm_b200 import _lib

    nbytes = int(lib.b200_lamport_workspace_bytes(world, max_numel))
    ws = [torch.empty(nbytes, dtype=torch.uint8, device=dev) for _ in range(world)]
    st = torch.cuda.current_stream(dev).cuda_stream
    for w in ws:
        _lib.check(lib.b200_lamport_workspace_init(w.data_ptr(), nbytes, st), "lamport_workspace_init")
    return ws, nbytes, (C.c_uint64 * world)(*[w.data_ptr() for w in ws])


@pytest.mark.parametrize("world", [2, 4, 8])
def test_emulated_lamport_all_reduce(cuda_lib, world):
    """Barrier-free all-reduce (one-shot up to 4 ranks, two-shot at 8), emulated ranks on one GPU: 10 calls in a row
    (every receive-buffer generation is reused three times) with CHANGING message shapes, words that equal the
    sentinel pattern in the inputs, exact against the fp32 sum rounded once."""
    from vllm_b200 import _lib

    dev = torch.device("cuda", 0)
    H = 1024
    ws, nbytes, tab = _lamport_ws(world, 64 * H, dev, cuda_lib)
    st = torch.cuda.current_stream(dev).cuda_stream
    g = torch.Generator(device=dev).manual_seed(3)
    for it, B in enumerate([48, 64, 7, 64, 1, 33, 64, 64, 5, 48]):
        xs = torch.randn(world, B, H, device=dev, generator=g).to(torch.bfloat16)
        xs[:, 0, :16] = -0.0   # sentinel words (pairs of -0) in every rank's message
        xs[0, B - 1, 8:24] = -0.0
        ref = xs.float().sum(0).to(torch.bfloat16)
        out = torch.full((world, B, H), float("nan"), dtype=torch.bfloat16, device=dev)
        rc = cuda_lib.b200_lamport_all_reduce(out.data_ptr(), xs.data_ptr(), tab, nbytes, -1, world, B, H, 2, st)
        _lib.check(rc, "lamport_all_reduce(emu)")
        torch.cuda.synchronize()
        for r in range(world):
            assert torch.equal(out[r], ref), (it, B, r)   # -0 == +0 under torch.equal: sign of zero is the documented caveat
    # a message that does not fit the workspace is rejected before launch
    big = torch.zeros(world, 65, H, dtype=torch.bfloat16, device=dev)
    rc = cuda_lib.b200_lamport_all_reduce(big.data_ptr(), big.data_ptr(), tab, nbytes, -1, world, 65, H, 2, st)
    assert rc != 0


@pytest.mark.parametrize("world", [2, 4, 8])
def test_emulated_lamport_all_reduce_rms_norm_quant(cuda_lib, world):
    """Fused Lamport all-reduce + residual add + RMSNorm + per-token FP8 quant == all-reduce then the stand-alone
    fused norm/quant kernel, bit for bit; more rows than CTAs (two rows per CTA) in some calls."""
    from vllm_b200 import _lib, ops

    dev = torch.device("cuda", 0)
    H = 2048
    ws, nbytes, tab = _lamport_ws(world, 80 * H, dev, cuda_lib)
    st = torch.cuda.current_stream(dev).cuda_stream
    g = torch.Generator(device=dev).manual_seed(9)
    w = (torch.rand(H, device=dev, generator=g) + 0.5).to(torch.bfloat16)
    for it, B in enumerate([40, 80, 3, 80, 40, 17, 40]):
        xs = torch.randn(world, B, H, device=dev, generator=g).to(torch.bfloat16)
        res = torch.randn(B, H, device=dev, generator=g).to(torch.bfloat16)
        summed = xs.float().sum(0).to(torch.bfloat16)
        res_ref = res.clone()
        xq_ref, xs_ref = ops.rms_norm_dynamic_per_token_quant(summed, w, 1e-5, residual=res_ref)
        xq = torch.empty(world, B, H, dtype=torch.float8_e4m3fn, device=dev)
        sc = torch.empty(world, B, dtype=torch.float32, device=dev)
        resid = res.unsqueeze(0).repeat(world, 1, 1).contiguous()
        rc = cuda_lib.b200_lamport_all_reduce_rms_norm_quant(xq.data_ptr(), sc.data_ptr(), xs.data_ptr(), tab, nbytes, -1,
                                                             world, w.data_ptr(), resid.data_ptr(), 1e-5, B, H, 2, st)
        _lib.check(rc, "lamport_all_reduce_rms_norm_quant(emu)")
        torch.cuda.synchronize()
        for r in range(world):
            assert torch.equal(xq[r].view(torch.uint8), xq_ref.view(torch.uint8)), (it, B, r)
            assert torch.equal(sc[r], xs_ref.reshape(-1))
            assert torch.equal(resid[r], res_ref)


@pytest.mark.parametrize("world,H", [(2, 1024), (8, 1024), (2, 8192), (8, 8192)])
def test_emulated_lamport_all_reduce_rms_norm_nvfp4_quant(cuda_lib, world, H):
    """Fused Lamport all-reduce + residual add + RMSNorm + NVFP4 quant == all-reduce, then b200_rms_norm_nvfp4_quant,
    bit for bit (codes, swizzled scales incl. the zeroed padding, residual)."""
    from vllm_b200 import _lib, ops

    dev = torch.device("cuda", 0)
    ws, nbytes, tab = _lamport_ws(world, 140 * H, dev, cuda_lib)
    st = torch.cuda.current_stream(dev).cuda_stream
    g = torch.Generator(device=dev).manual_seed(21)
    w = (torch.rand(H, device=dev, generator=g) + 0.5).to(torch.bfloat16)
    gs = torch.tensor([448.0 * 6.0 / 8.0], device=dev)
    for it, B in enumerate([40, 130, 3, 140]):
        xs = torch.randn(world, B, H, device=dev, generator=g).to(torch.bfloat16)
        res = torch.randn(B, H, device=dev, generator=g).to(torch.bfloat16)
        summed = xs.float().sum(0).to(torch.bfloat16)
        res_ref = res.clone()
        q_ref, sf_ref = ops.rms_norm_nvfp4_quant(summed, w, 1e-5, gs, residual=res_ref)
        rows_pad, sf_cols = (B + 127) // 128 * 128, (H // 16 + 3) // 4 * 4
        q = torch.empty(world, B, H // 2, dtype=torch.uint8, device=dev)
        sf = torch.full((world, rows_pad, sf_cols), 0x55, dtype=torch.uint8, device=dev)
        resid = res.unsqueeze(0).repeat(world, 1, 1).contiguous()
        rc = cuda_lib.b200_lamport_all_reduce_rms_norm_nvfp4_quant(
            q.data_ptr(), sf.data_ptr(), xs.data_ptr(), tab, nbytes, -1, world, w.data_ptr(), resid.data_ptr(),
            gs.data_ptr(), 1e-5, B, H, 2, st)
        _lib.check(rc, "lamport_all_reduce_rms_norm_nvfp4_quant(emu)")
        torch.cuda.synchronize()
        for r in range(world):
            assert torch.equal(q[r], q_ref), (it, B, r)
            assert torch.equal(sf[r].flatten(), sf_ref.view(torch.uint8).flatten()), (it, B, r)
            assert torch.equal(resid[r], res_ref)


# ------------------------------------------------------------------------------- real ranks
def _worker(rank, world, port, q, mode):
    nvls = mode == "nvls"
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), B200_NVLS="force" if nvls else "0",
                      B200_AR_LAMPORT="1" if mode == "lamport" else "0")
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from vllm_b200 import ops
    from vllm_b200.peer_comm import PeerAllReduce

    B, H = 48, 1024
    par = PeerAllReduce(dist.group.WORLD, dev, 2, B * H)
    used_nvls = bool(par.multicast_ptr)
    assert par.lamport == (mode == "lamport")
    if nvls and not used_nvls and rank == 0:
        print("NVLS requested but the fabric exposes no multicast mapping: peer-load path tested instead", flush=True)
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
        # fp32 accumulation + one final rounding.  NCCL's fp32 sum order may differ from rank order by an ulp
        # of fp32, which can flip a bf16 rounding at most in isolated elements: compare within 1 bf16 ulp and
        # require exact agreement BETWEEN the ranks (the property the decode step relies on).
        ok = ok and torch.allclose(got.float(), ref, rtol=8e-3, atol=1e-6)
        allg = [torch.empty_like(got) for _ in range(world)]
        dist.all_gather(allg, got.contiguous())
        ok = ok and all(torch.equal(allg[0], t) for t in allg)
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
        q.put((ok, used_nvls))
    torch.cuda.synchronize()
    dist.barrier()
    os._exit(0)


@pytest.mark.parametrize("mode", ["lamport", "nvls", "peer"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_peer_all_reduce_real_ranks(world, mode):
    """lamport: the barrier-free protocol (default); nvls: multimem.ld_reduce / multimem.st (in-switch sum) when
    torch's symmetric memory exposes a multicast mapping; peer: the flag-barrier peer-load kernels."""
    nvls = mode
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q, mode)) for r in range(world)]
    for p in procs:
        p.start()
    ok, used_nvls = q.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
    print(f"world {world}: mode {mode}, nvls used {used_nvls}")
    assert ok
