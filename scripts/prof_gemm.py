#!/usr/bin/env python
"""Run one FP8 GEMM shape a few times (for ncu): python scripts/prof_gemm.py --m 128 --n 4096 --k 4096 [--nvfp4]"""
import argparse, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vllm_b200 import ops

ap = argparse.ArgumentParser()
ap.add_argument("--m", type=int, default=128); ap.add_argument("--n", type=int, default=4096)
ap.add_argument("--k", type=int, default=4096); ap.add_argument("--reps", type=int, default=6)
ap.add_argument("--nvfp4", action="store_true")
ap.add_argument("--trace", action="store_true", help="print the decode kernel's per-phase %%globaltimer stamps")
ap.add_argument("--mode", type=int, default=-1, help="b200_debug_gemm_mode (4 = 2-CTA kernel for M <= 256)")
a = ap.parse_args()
dev = "cuda:0"
sets = 6
if a.mode != -1:
    from vllm_b200 import _lib as _l
    _l.load().b200_debug_gemm_mode(a.mode)
if a.nvfp4:
    gs = torch.tensor([1.0], device=dev)
    W = [ops.scaled_fp4_quant((torch.randn(a.n, a.k, device=dev) * 0.5).to(torch.bfloat16), gs) for _ in range(sets)]
    X = [ops.scaled_fp4_quant(torch.randn(a.m, a.k, device=dev).to(torch.bfloat16), gs) for _ in range(sets)]
    f = lambda i: ops.cutlass_scaled_fp4_mm(X[i][0], W[i][0], X[i][1], W[i][1], gs, torch.bfloat16)
else:
    W = [(torch.randn(a.n, a.k, device=dev) * 0.05).to(torch.float8_e4m3fn) for _ in range(sets)]
    X = [torch.randn(a.m, a.k, device=dev).to(torch.float8_e4m3fn) for _ in range(sets)]
    sa, sb = torch.rand(a.m, 1, device=dev) + 0.5, torch.rand(1, device=dev) + 0.5
    f = lambda i: ops.cutlass_scaled_mm(X[i], W[i].t(), sa, sb, torch.bfloat16)
for i in range(a.reps):
    y = f(i % sets)
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
if a.trace:
    from vllm_b200 import _lib
    names = ["entry", "setup", "ring_filled", "first_full", "last_commit", "acc_ready", "published", "gathered",
             "out_done", "epi_end", "exit"]
    if a.mode == 4:  # the 2-CTA kernel's stamps
        names = ["entry", "setup", "weights_issued", "first_full", "last_commit", "acc_ready", "pdl_wait_done",
                 "kb_STAGES_full", "out_done", "-", "exit"]
    buf = torch.zeros(512, 16, dtype=torch.int64, device=dev)
    lib = _lib.load()
    for rep in range(3):
        buf.zero_()
        lib.b200_debug_gemm_trace(buf.data_ptr())
        f(rep % sets)
        torch.cuda.synchronize()
        lib.b200_debug_gemm_trace(None)
        t = buf.cpu()
        used = t[:, 0] > 0
        t = t[used].double()
        t0 = t[:, 0].min()
        print(f"--- rep {rep}: {int(used.sum())} CTAs, span {float(t[:, 10].max() - t0) / 1e3:.2f} us "
              "(stamps relative to the first CTA's entry, us: min / median / max over CTAs)")
        for i, nm in enumerate(names):
            col = t[:, i]
            col = col[col > 0]
            if len(col) == 0:
                continue
            r = (col - t0) / 1e3
            print(f"  {nm:12s} {float(r.min()):7.2f} {float(r.median()):7.2f} {float(r.max()):7.2f}   n={len(col)}")
