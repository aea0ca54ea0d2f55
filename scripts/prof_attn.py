#!/usr/bin/env python
"""Run the fp8 paged-attention decode kernel at the BASELINE cfg[2] shape (batch 128 x ctx 4K, 32/8 heads)
and print its per-phase %globaltimer stamps and per-barrier wait shares:
    python scripts/prof_attn.py [--splits N] [--trace]"""
import argparse, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vllm_b200 import ops, _lib

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=128); ap.add_argument("--ctx", type=int, default=4096)
ap.add_argument("--hq", type=int, default=32); ap.add_argument("--hkv", type=int, default=8)
ap.add_argument("--splits", type=int, default=0); ap.add_argument("--trace", action="store_true")
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()
DEV, D, bs = "cuda:0", 128, 16
B, ctx, Hq, Hkv = a.batch, a.ctx, a.hq, a.hkv
bps = ctx // bs
nb = B * bps + 1
kv = torch.randint(0, 256, (2, nb, bs, Hkv, D), device=DEV, dtype=torch.uint8) & 0x77
bt = (torch.randperm(nb - 1, device=DEV)[: B * bps] + 1).reshape(B, bps).to(torch.int32)
sl = torch.full((B,), ctx, device=DEV, dtype=torch.int32)
q = torch.randn(B, Hq, D, device=DEV).to(torch.bfloat16)
out = torch.empty_like(q)
sc = torch.tensor([0.05], device=DEV)
splits = a.splits or ops.paged_decode_num_splits(B, Hq, Hkv, ctx, DEV, D, "fp8")
ws = torch.empty(max(ops.paged_decode_workspace_bytes(B, Hq, D, splits), 16), dtype=torch.uint8, device=DEV)
f = lambda: ops.paged_attention_decode(out, q, kv[0], kv[1], bt, sl, D ** -0.5, ctx, "fp8", sc, sc,
                                       num_splits=splits, workspace=ws)
for _ in range(a.reps):
    f()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); f(); e1.record(); torch.cuda.synchronize()
print(f"splits {splits}  one launch: {e0.elapsed_time(e1) * 1e3:.1f} us")
if a.trace:
    names = ["entry", "setup", "tma_issue", "full", "qk_issued", "s_full", "s_loaded", "max_done", "p_ready",
             "pv_start", "pv_issued", "o_full", "folded", "loop_end", "exit"]
    buf = torch.zeros(4096, 32, dtype=torch.int64, device=DEV)
    lib = _lib.load()
    lib.b200_debug_attn_trace(buf.data_ptr())
    f(); torch.cuda.synchronize()
    lib.b200_debug_attn_trace(None)
    t = buf.cpu().double()
    t = t[t[:, 0] > 0]
    print(f"{len(t)} CTAs traced; per-CTA stamps relative to the CTA's own entry (us): min / median / max")
    for i, nm in enumerate(names):
        ok = t[:, i] > 0
        if ok.sum() == 0:
            continue
        r = (t[ok, i] - t[ok, 0]) / 1e3
        print(f"  {nm:10s} {float(r.min()):7.2f} {float(r.median()):7.2f} {float(r.max()):7.2f}   n={int(ok.sum())}")
    wn = ["k_empty(K producer)", "v_empty(V producer)", "k_full(QK issuer)", "s_empty(QK issuer)", "v_full(PV issuer)",
          "p_full(PV issuer)", "o_empty(PV issuer)", "s_full(softmax g0)", "o_full(softmax g0)", "group barrier(g0)"]
    tot = t[:, 26]
    print("  share of the softmax loop time each role spends blocked at a barrier (median over CTAs):")
    for i, nm in enumerate(wn):
        print(f"    {nm:22s} {float((t[:, 16 + i] / tot).median()) * 100:5.1f} %")
    print(f"    K producer: TMA issue section {float((t[:, 27] / t[:, 28]).median()) * 100:.1f} % of its loop, "
          f"loop {float(t[:, 28].median()):.0f} cycles vs softmax loop {float(tot.median()):.0f} cycles")
    span = (t[:, 14].max() - t[:, 0].min()) / 1e3
    print(f"  kernel span {float(span):.1f} us; CTA lifetime median {float(((t[:, 14] - t[:, 0]) / 1e3).median()):.2f} us")
