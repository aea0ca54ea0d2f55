#!/usr/bin/env python
"""Verifies the TMEM data-path layout of tcgen05.mma.cta_group::2 with UMMA M = 128 (64 rows per CTA), which the
decode-regime 2-CTA FP8 kernel relies on: runs ONE 128 x 128 tile with unit scales, dumps the raw TMEM image of both
CTAs (b200_debug_gemm_dump) and matches every (cta, lane, column) against the fp32 reference product."""
import os, sys
from pathlib import Path
os.environ["B200_GEMM_2SM_BN"] = "128"
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vllm_b200 import _lib, ops

dev = torch.device("cuda:0")
M, N, K = 128, 128, 256
g = torch.Generator().manual_seed(3)
a = torch.randn(M, K, generator=g).to(torch.float8_e4m3fn).to(dev)
b = torch.randn(N, K, generator=g).to(torch.float8_e4m3fn).to(dev)
one = torch.ones(1, device=dev)
ref = a.float() @ b.float().t()
lib = _lib.load()
dump = torch.full((2, 128, 64), float("nan"), device=dev)
lib.b200_debug_gemm_dump(dump.data_ptr())
lib.b200_debug_gemm_mode(4)
out = ops.cutlass_scaled_mm(a, b.t(), one, one, torch.bfloat16)
torch.cuda.synchronize()
lib.b200_debug_gemm_dump(None)
lib.b200_debug_gemm_mode(-1)
err = (out.float() - ref).abs().max().item()
print("max |out - ref| =", err, "(bf16 rounding of |ref| <=", ref.abs().max().item(), ")")
d = dump.cpu()
r = ref.cpu()
# match each dumped value to its (m, n)
flat = r.flatten()
vals, order = flat.sort()
for cta in range(2):
    for lane in (0, 1, 31, 32, 63, 64, 65, 96, 127):
        row = d[cta, lane]
        hits = []
        for col in (0, 1, 31, 32, 63):
            v = row[col].item()
            if v != v:
                hits.append((col, "nan"))
                continue
            i = torch.searchsorted(vals, torch.tensor(v)).clamp(0, vals.numel() - 1)
            best = min((abs(vals[j].item() - v), j) for j in range(max(0, i - 2), min(vals.numel(), i + 3)))
            idx = order[best[1]].item()
            hits.append((col, (idx // N, idx % N), round(best[0], 4)))
        print(f"cta {cta} lane {lane:3d}:", hits)
ok = err < 0.51 * 2 ** -7 * ref.abs().max().item() + 1e-3
print("LAYOUT_OK" if ok else "LAYOUT_MISMATCH")
