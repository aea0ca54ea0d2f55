#!/usr/bin/env python
"""Decode-step benchmark of the B200 hot path (see DESIGN.md "Measurement").

  python bench.py --gpus N --steps K --warmup W            # this repo (CUDA, C ABI)
  python bench.py --impl reference --gpus N ...            # reference CPU path, host cores

A step = one decode step (one new token for every sequence of the batch) of a
Llama-3.1-8B-shaped model with FP8 W8A8 linears + FP8 KV cache, batch 128 x ctx 4096
(BASELINE.json configs[2], the config the metric "decode tokens/s/GPU Llama-3.1-8B FP8 at
TP1-8" is quoted on), tensor-parallel over N GPUs.  Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOAD = "configs[2]: Llama-3.1-8B FP8 W8A8 per-tensor weights + FP8 KV cache, batch 128 x ctx 4096, TP=N"
METRIC = "decode tokens/s (whole job; /n_gpus = per GPU), Llama-3.1-8B FP8 W8A8 + FP8 KV, batch 128 x ctx 4096"


def peaks() -> tuple[dict, str]:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return json.loads(p.read_text()), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx[0] if mx else None, "reasons": reasons,
                "samples": len(sm)}


# ------------------------------------------------------------------------ reference arm
def cpu_sample(steps: int, warmup: int, threads: int | None = None) -> dict:
    """The reference's own CPU implementation of the path, bounded sample: ONE decoder layer's hot path
    (KV-cache write + paged-attention decode through the reference's compiled CPU kernels when
    oracle/_ref is usable, else the oracle port; the 4 linears as torch CPU matmuls on bf16 weights —
    the reference CPU backend's FP8 GEMM needs oneDNN which cannot be built offline) for 8 sequences of
    ctx 4096; tokens/s = 8 / (32 layers x t_layer)."""
    from oracle import oracle as O
    from oracle import ref_cpu

    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    Bs, ctx, Hq, Hkv, D, H, I, L = 8, 4096, 32, 8, 128, 4096, 14336, 32
    use_ref = ref_cpu.available()
    bs = 32 if use_ref else 16
    g = torch.Generator().manual_seed(0)
    nb = Bs * (ctx // bs) + 1
    dt = torch.bfloat16
    q = torch.randn(Bs, Hq, D, generator=g).to(dt)
    k_new = torch.randn(Bs, Hkv, D, generator=g).to(dt)
    v_new = torch.randn(Bs, Hkv, D, generator=g).to(dt)
    bt = (torch.randperm(nb - 1, generator=g)[: Bs * (ctx // bs)] + 1).reshape(Bs, -1).to(torch.int32)
    sl = torch.full((Bs,), ctx, dtype=torch.int32)
    slots = bt[:, -1].long() * bs + (ctx - 1) % bs
    x = torch.randn(Bs, H, generator=g).to(dt)
    wqkv = torch.randn((Hq + 2 * Hkv) * D, H, generator=g).to(dt)
    wo = torch.randn(H, Hq * D, generator=g).to(dt)
    wgu = torch.randn(2 * I, H, generator=g).to(dt)
    wd = torch.randn(H, I, generator=g).to(dt)
    if use_ref:
        isa = ref_cpu.pick_isa(dt, bs, D)
        kc = ((torch.rand(nb, Hkv, bs, D, generator=g) * 2 - 1) * D ** -0.5).to(dt)
        vc = ((torch.rand(nb, Hkv, bs, D, generator=g) * 2 - 1) * D ** -0.5).to(dt)
    else:
        kc = ((torch.rand(nb, bs, Hkv, D, generator=g) * 2 - 1) * D ** -0.5).to(dt)
        vc = ((torch.rand(nb, bs, Hkv, D, generator=g) * 2 - 1) * D ** -0.5).to(dt)

    def layer():
        qkv = x @ wqkv.t()
        if use_ref:
            ref_cpu.reshape_and_cache(k_new, v_new, kc, vc, slots, isa)
            a = ref_cpu.decode_attention(q, kc, vc, bt, sl, D ** -0.5, isa)
        else:
            O.reshape_and_cache_flash(k_new, v_new, kc, vc, slots, "auto", torch.ones(1), torch.ones(1))
            a = O.paged_attention_decode(q, kc, vc, bt, sl, D ** -0.5).to(dt)
        h = a.reshape(Bs, -1) @ wo.t()
        gu = h @ wgu.t()
        act = torch.nn.functional.silu(gu[:, :I]) * gu[:, I:]
        return act @ wd.t() + qkv[:, :1]

    for _ in range(max(1, warmup)):
        layer()
    t0 = time.perf_counter()
    for _ in range(steps):
        layer()
    dt_s = (time.perf_counter() - t0) / steps
    return {"value": Bs / (L * dt_s), "unit": "tokens/s", "cores": threads, "kind": "reference" if use_ref else "port",
            "sample": f"1 decoder layer (cache write + paged attention via "
                      f"{'reference csrc/cpu/cpu_attn.cpp kernels' if use_ref else 'oracle port'}, 4 bf16 linears via torch CPU) "
                      f"x 8 seqs x ctx 4096, scaled to 32 layers; {dt_s * 1e3:.1f} ms/layer",
            "ms_per_layer": dt_s * 1e3}


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = min(args.steps, 20)
    res = cpu_sample(steps, min(args.warmup, 3))
    line = {"impl": "reference", "metric": METRIC, "value": res["value"], "unit": "tokens/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": min(args.warmup, 3), "ms_per_step": 32 * res["ms_per_layer"] * (128 / 8),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "note": "CPU arm: bounded sample, see cpu_baseline.sample"},
            "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": res["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def attn_traffic(world: int, B: int, ctx: int) -> tuple[float | None, str]:
    """Per-launch DRAM bytes of the dominant kernel from the committed ncu capture (profiles/attn_traffic.json,
    written by scripts/ncu_summary.py traffic) - only while the kernel sources are still the ones that were
    profiled (digest match) and the workload is the profiled one; otherwise null."""
    import hashlib

    f = ROOT / "profiles" / "attn_traffic.json"
    if not f.exists():
        return None, "no profiles/attn_traffic.json"
    rec = json.loads(f.read_text())
    h = hashlib.sha256()
    for name in ("paged_decode.cu", "paged_decode.cuh", "common.cuh"):
        h.update((ROOT / "vllm_b200" / "csrc" / name).read_bytes())
    if h.hexdigest()[:16] != rec.get("kernel_src_sha16"):
        return None, "profiles/attn_traffic.json was captured from different kernel sources (digest mismatch)"
    if not (world == 1 and B == 128 and ctx == 4096):
        return None, "capture is for configs[2] TP1 only"
    return rec["dram_bytes_per_launch"], (f"dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full, "
                                          f"profiles/attn_traffic.json (kernel sources {rec['kernel_src_sha16']})")


# ------------------------------------------------------------------------------ our arm
def run_ours(args) -> None:
    import torch.distributed as dist

    from vllm_b200 import _lib, ops
    from vllm_b200.decode_runner import LLAMA_8B, DecodeRunner

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: vllm_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    _lib.load()

    def stage(msg: str) -> None:  # progress on stderr (stdout carries exactly one JSON line)
        if world > 1 or os.environ.get("B200_BENCH_VERBOSE"):
            print(f"[bench rank {rank}/{world}] {msg}", file=sys.stderr, flush=True)

    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD
        stage("process group up")
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    from vllm_b200.decode_runner import LLAMA_70B

    # cfg3 (default, the headline): BASELINE configs[2].  cfg4: configs[3] - Llama-3.1-70B shapes, NVFP4 block-scaled
    # weights + fp8 KV, batch 256 x ctx 8192, meant for TP=8 (43 GB of KV + 4.3 GB of weights per rank).
    cfg4 = args.config == "cfg4"
    model, wq = (LLAMA_70B, "nvfp4") if cfg4 else (LLAMA_8B, "fp8")
    if cfg4 and args.batch == 128 and args.ctx == 4096:  # the config's own shape unless overridden
        args.batch, args.ctx = 256, 8192
    workload = (f"configs[3]: Llama-3.1-70B NVFP4 block-scaled weights + FP8 KV cache, batch {args.batch} x ctx {args.ctx}, "
                f"TP={world}") if cfg4 else WORKLOAD.replace("TP=N", f"TP={world}")
    metric = (f"decode tokens/s (whole job), Llama-3.1-70B NVFP4 + FP8 KV, batch {args.batch} x ctx {args.ctx}"
              if cfg4 else METRIC)
    B, ctx = args.batch, args.ctx
    emu = args.emulate_tp if (world == 1 and args.emulate_tp > 1) else 0
    runner = DecodeRunner(model, B, ctx, kv_cache_dtype="fp8", weight_quant=wq, tp_size=emu or world, tp_rank=rank,
                          device=dev, tp_group=group, num_layers=args.layers, skip_collectives=bool(emu))
    if emu:
        workload += f" [INVALID as a result: rank 0's shard of TP={emu} on ONE GPU, all-reduces skipped - development aid]"
    runner.set_batch(ctx)
    stage("runner built")
    if args.profile_step:
        # for `ncu --profile-from-start off`: exactly one eager step between profiler start/stop
        runner.step_eager()
        runner.step_eager()
        torch.cuda.synchronize(dev)
        torch.cuda.profiler.start()
        runner.step_eager()
        torch.cuda.synchronize(dev)
        torch.cuda.profiler.stop()
        print(json.dumps({"profile_step": True, "launches_ours": _lib.num_kernels_launched()}))
        return
    runner.capture(warmup=2)
    stage("step captured")

    def barrier():
        if world > 1:
            dist.barrier(group=group)
        torch.cuda.synchronize(dev)

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        return float(t.item())

    # ---- device-resident timing: K graph replays
    for _ in range(max(args.warmup, 3)):
        runner.step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        runner.step()
    e1.record()
    barrier()
    ms_dev = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    stage(f"device-resident timing done: {ms_dev:.3f} ms/step")
    clocks = sampler.stop() if rank == 0 else None

    # ---- end to end: pinned host inputs -> H2D -> step -> D2H of the sampled tokens, every step
    h_ids = torch.randint(0, model.vocab, (B,), dtype=torch.int64).pin_memory()
    h_pos = torch.full((B,), ctx - 1, dtype=torch.int64).pin_memory()
    h_slots = runner.slot_mapping.cpu().pin_memory()
    h_lens = runner.seq_lens.cpu().pin_memory()
    h_bt = runner.block_table.cpu().pin_memory()
    h_out = torch.empty(B, dtype=torch.int64).pin_memory()
    h2d = sum(t.numel() * t.element_size() for t in (h_ids, h_pos, h_slots, h_lens, h_bt))
    d2h = h_out.numel() * h_out.element_size()

    def e2e_step():
        runner.input_ids.copy_(h_ids, non_blocking=True)
        runner.positions.copy_(h_pos, non_blocking=True)
        runner.slot_mapping.copy_(h_slots, non_blocking=True)
        runner.seq_lens.copy_(h_lens, non_blocking=True)
        runner.block_table.copy_(h_bt, non_blocking=True)
        runner.step()
        h_out.copy_(runner.next_tokens, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()  # the caller consumes the tokens each step

    for _ in range(3):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    # the caller-visible time: host wall clock around the loop (it includes the per-step stream synchronise);
    # the device-side event span is reported beside it and can only be shorter
    ms_e2e_dev = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    ms_e2e = max(max_over_ranks(wall_ms) / args.steps, ms_e2e_dev)

    # ---- dominant kernel (paged-attention decode) timed alone, CUDA events on its launch stream,
    #      cycling through the layers' distinct 1 GB caches so no launch finds its KV in L2
    attn_out = torch.empty(B, runner.Hq, runner.D, device=dev, dtype=torch.bfloat16)
    q = torch.randn(B, runner.Hq, runner.D, device=dev).to(torch.bfloat16)

    def attn(layer):
        kc, vc = runner.kv_caches[layer].unbind(0)
        ops.paged_attention_decode(attn_out, q, kc, vc, runner.block_table, runner.seq_lens, runner.scale, ctx, "fp8",
                                   runner.k_scale, runner.v_scale, num_splits=runner.num_splits, workspace=runner.attn_ws)

    for l in range(min(3, runner.L)):
        attn(l)
    torch.cuda.synchronize(dev)
    reps = max(1, 64 // runner.L) * runner.L
    e0.record()
    for i in range(reps):
        attn(i % runner.L)
    e1.record()
    torch.cuda.synchronize(dev)
    attn_ms = e0.elapsed_time(e1) / reps
    attn_bytes = runner.attention_bytes_per_layer(ctx)
    pk, pk_kind = peaks()
    achieved = attn_bytes / (attn_ms * 1e-3) / 1e9

    traffic, traffic_src = attn_traffic(world, B, ctx)
    if args.kernel_table:
        # every rank replays (the step contains collectives), rank 0 records: the launch list of the N-GPU step
        from torch.profiler import ProfilerActivity, profile

        barrier()
        if rank == 0:
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                for _ in range(5):
                    runner.step()
                torch.cuda.synchronize(dev)
            rows = sorted(((e.key, e.count, e.device_time_total) for e in prof.key_averages() if e.device_time_total > 0),
                          key=lambda r: -r[2])
            tot = sum(r[2] for r in rows)
            out = [f"# bench.py --gpus {world}: rank 0, 5 replayed steps, {tot / 5e3:.3f} ms of kernels per step "
                   f"(step {ms_dev:.3f} ms)", f"{'us/step':>9} {'n/step':>7} {'avg_us':>8} {'share':>6}  kernel"]
            for k, n, t in rows[:30]:
                out.append(f"{t / 5:9.1f} {n / 5:7.1f} {t / n:8.2f} {100 * t / tot:5.1f}%  {k[:140]}")
            Path(args.kernel_table).write_text("\n".join(out) + "\n")
        else:
            for _ in range(5):
                runner.step()
            torch.cuda.synchronize(dev)
        barrier()
    # ---- untimed self-check at N > 1: does the TP=N step compute the TP=1 function?  (small 2-layer model built from
    #      ONE seeded checkpoint and sharded through vllm_b200.tp_shard; every rank compares against its local TP=1 run)
    tp_check = None
    if world > 1 and not args.no_tp_check:
        from vllm_b200.decode_runner import tp_equivalence_check

        try:
            res = tp_equivalence_check(world, rank, group, dev)
        except Exception as e:  # noqa: BLE001 - reported, never fatal for the measurement
            res = {"hidden_rel_l2": float("nan"), "tokens_equal_frac": 0.0, "peer_allreduce": None, "error": repr(e)}
        allres = [None] * world
        dist.all_gather_object(allres, res, group=group)
        tp_check = {"errors": [r["error"] for r in allres if "error" in r],"hidden_rel_l2_max_over_ranks": max(r["hidden_rel_l2"] for r in allres),
                    "tokens_equal_frac_min_over_ranks": min(r["tokens_equal_frac"] for r in allres),
                    "peer_allreduce": res["peer_allreduce"], "bound_rel_l2": 0.10,
                    "what": "TP=N vs TP=1 hidden state after the final norm, 2-layer Llama-shaped model, batch 16 x ctx 500"}
    if rank == 0:
        tok_s = B / (ms_dev * 1e-3)
        step_bytes = runner.L * (attn_bytes + runner.weight_bytes_per_layer())
        line = {
            "metric": metric, "value": tok_s, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None,
            "dtype": ("nvfp4(e2m1 + e4m3 block scales) GEMM operands + fp8 KV, fp32 accumulate, bf16 activations" if cfg4
                      else "fp8(e4m3) GEMM operands + fp8 KV, fp32 accumulate, bf16 activations"),
            "data": "synthetic",
            "config": {"workload": workload, "batch": B, "ctx": ctx,
                       "layers": runner.L, "block_size": 16, "parallelism": f"tp{world}",
                       "l2": f"per-step working set (KV + weights) >> 126 MB L2; kernel loop cycles {runner.L} distinct caches",
                       "cuda_graph": True, "attn_splits": runner.num_splits,
                       "steady_state": "every timed step replays the SAME decode position (positions / seq_lens / "
                                       "slot_mapping are not advanced: the new token's KV slot is rewritten each replay); "
                                       f"per-step work is identical to a real step at ctx {ctx}"},
            "tokens_per_s_per_gpu": tok_s / world,
            "e2e": {"value": B / (ms_e2e * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e, "ms_per_step_device_events": ms_e2e_dev},
            "gpu_launches": runner.launches_per_step * args.steps,
            "gpu_launches_per_step": runner.launches_per_step,
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": "paged_decode_kernel<bf16,128,fp8_e4m3> (+combine)",
                         "achieved": achieved, "peak": pk["hbm_gbs"], "peak_kind": f"{pk_kind} copy bandwidth (burst)",
                         "unit": "GB/s", "frac": achieved / pk["hbm_gbs"], "frac_of_8TBps_nameplate": achieved / 8000.0,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "algorithmic_bytes_per_launch": attn_bytes, "ms_per_launch": attn_ms},
            "step_model": {"algorithmic_bytes_per_step": step_bytes, "hbm_frac_of_step": step_bytes / (ms_dev * 1e-3) / 1e9 / pk["hbm_gbs"]},
        }
        if tp_check is not None:
            line["tp_check"] = tp_check
        if world == 1 and not args.no_cpu_baseline and not cfg4:
            try:
                res = cpu_sample(3, 1)
                line["cpu_baseline"] = {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")}
            except Exception as e:  # the baseline is reported, never required
                line["cpu_baseline"] = {"value": None, "unit": "tokens/s", "cores": os.cpu_count(), "kind": "port",
                                        "sample": f"failed: {e!r}"}
        print(json.dumps(line))
    if world > 1:
        # Tear-down: a captured graph that holds NCCL kernels makes ncclCommDestroy block (observed: the JSON line
        # printed, then both ranks hung in destroy_process_group until the launcher's timeout).  Release the
        # graph, drain the device, agree that every rank is done, and leave without the communicator teardown.
        sys.stdout.flush()
        runner.graph = None
        torch.cuda.synchronize(dev)
        dist.barrier(group=group)
        torch.cuda.synchronize(dev)
        stage("done")
        sys.stderr.flush()
        os._exit(0)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--ctx", type=int, default=4096)
    ap.add_argument("--layers", type=int, default=None, help="debug only: fewer layers (INVALID as a bench number)")
    ap.add_argument("--config", default="cfg3", choices=["cfg3", "cfg4"],
                    help="cfg3 = BASELINE configs[2] (default, the headline); cfg4 = configs[3] (70B NVFP4, for --gpus 8)")
    ap.add_argument("--emulate-tp", type=int, default=0, help="debug only (INVALID as a bench number): with --gpus 1, run "
                    "rank 0's shard of a TP=N model with the all-reduces skipped")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-tp-check", action="store_true")
    ap.add_argument("--kernel-table", default="", help="file: rank 0 writes the per-kernel device time of 5 replayed "
                                                       "steps (torch profiler / CUPTI; run AFTER the timed regions)")
    ap.add_argument("--profile-step", action="store_true", help="run one eager step between cudaProfilerStart/Stop")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
