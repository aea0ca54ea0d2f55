#!/usr/bin/env python
"""Engine-level decode throughput of the INSTALLED, UNMODIFIED vLLM with the B200 plugin OFF vs ON (GPU box only).

  python scripts/vllm_bench.py [--batch 128] [--prompt-len 64] [--decode 4032] [--arms stock,b200]

Model: Llama-3.1-8B architecture, dummy weights, online FP8 W8A8 (`quantization="fp8"`) + fp8 KV cache, i.e. the
numerics of BASELINE configs[2].  Workload: `batch` requests with a SHORT prompt and a LONG greedy decode, so that
nearly all time is spent in decode steps whose context grows from prompt_len to prompt_len + decode (default 64 ->
4096; the mean context is ~2K, NOT the fixed 4K of bench.py - stated in the output).  A short prompt on purpose: our
backend serves prefill tokens as single-query rows (correct, O(n^2) KV reads), prefill is outside
the path this plugin accelerates.

Each arm runs in its own subprocess (one engine per process); which kernels were selected is proved with the
engine's own log lines ("Using ... attention backend", "Selected ... for ...").  One JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent

LLAMA_8B = {
    "architectures": ["LlamaForCausalLM"], "model_type": "llama", "hidden_size": 4096, "intermediate_size": 14336,
    "num_hidden_layers": 32, "num_attention_heads": 32, "num_key_value_heads": 8, "head_dim": 128, "vocab_size": 128256,
    "max_position_embeddings": 8192, "rms_norm_eps": 1e-5, "rope_theta": 500000.0, "hidden_act": "silu",
    "tie_word_embeddings": False, "torch_dtype": "bfloat16", "bos_token_id": 1, "eos_token_id": 2,
    "attention_bias": False, "mlp_bias": False,
}


def worker(a) -> None:
    sys.path.insert(0, str(ROOT))
    os.environ.setdefault("VLLM_ENABLE_V1_MULTIPROCESSING", "0")
    kind = a.arm.split(":", 1)[0]
    backend = a.arm.split(":", 1)[1] if ":" in a.arm else None
    if kind.startswith("b200"):
        from vllm_b200 import plugin

        plugin.register()
        backend = "CUSTOM"
    else:
        os.environ["B200_VLLM_LINEAR"] = "0"
        os.environ["VLLM_DISABLED_KERNELS"] = "B200FP8ScaledMMLinearKernel,B200NvFp4LinearKernel"
    import torch
    from vllm import LLM, SamplingParams

    kw = dict(model=a.dir, load_format="dummy", skip_tokenizer_init=True, dtype="bfloat16",
              max_model_len=a.prompt_len + a.decode + 16, gpu_memory_utilization=0.6, seed=0, block_size=16,
              max_num_seqs=a.batch, enable_prefix_caching=False, quantization="fp8", kv_cache_dtype="fp8")
    if backend:
        kw["attention_backend"] = backend
    llm = LLM(**kw)
    g = torch.Generator().manual_seed(0)
    prompts = [{"prompt_token_ids": torch.randint(10, 100000, (a.prompt_len,), generator=g).tolist()} for _ in range(a.batch)]

    def run(n):
        sp = SamplingParams(temperature=0.0, max_tokens=n, ignore_eos=True, detokenize=False)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        outs = llm.generate(prompts, sp, use_tqdm=False)
        torch.cuda.synchronize()
        return time.perf_counter() - t0, sum(len(o.outputs[0].token_ids) for o in outs)

    run(8)  # warm-up: graphs captured, lazy modules loaded
    if a.profile:
        # per-kernel device time of ~48 decode steps (CUPTI sees the kernels inside replayed CUDA graphs)
        from torch.profiler import ProfilerActivity, profile

        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            run(48)
        rows = sorted(((e.key, e.count, e.device_time_total) for e in prof.key_averages() if e.device_time_total > 0),
                      key=lambda r: -r[2])
        tot = sum(r[2] for r in rows)
        lines = [f"# arm {a.arm}: {tot / 1e3:.1f} ms of kernels over prefill + 48 decode steps of batch {a.batch}",
                 f"{'total_ms':>9} {'n':>6} {'avg_us':>8} {'share':>6}  kernel"]
        for k, n, t in rows[:40]:
            lines.append(f"{t / 1e3:9.2f} {n:6d} {t / n:8.2f} {100 * t / tot:5.1f}%  {k[:150]}")
        Path(a.profile).write_text("\n".join(lines) + "\n")
    t_short, n_short = run(16)          # prefill + 16 decode steps
    t_long, n_long = run(a.decode)      # prefill + `decode` steps
    # decode-only rate: the two runs share the same prefill, the difference is (decode - 16) steps of all requests
    rate = (n_long - n_short) / (t_long - t_short)
    sel = {"b200_lib_loaded": False, "b200_launches": 0}
    try:
        import vllm_b200._lib as L

        sel = {"b200_lib_loaded": L.is_loaded(), "b200_launches": L.num_kernels_launched() if L.is_loaded() else 0}
    except Exception:
        pass
    Path(a.out).write_text(json.dumps({"decode_tokens_per_s": rate, "t_long_s": t_long, "t_short_s": t_short,
                                       "tokens_long": n_long, "ms_per_decode_step": 1e3 * a.batch / rate, "sel": sel}))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--prompt-len", type=int, default=64)
    ap.add_argument("--decode", type=int, default=4032)
    ap.add_argument("--arms", default="stock,b200")
    ap.add_argument("--timeout", type=int, default=1500)
    ap.add_argument("--profile", default="", help="directory: also write a per-kernel device-time table per arm")
    ap.add_argument("--log-dir", default="", help="directory: keep every arm's full engine log")
    ap.add_argument("--worker", action="store_true")
    ap.add_argument("--arm", default="stock")
    ap.add_argument("--dir")
    ap.add_argument("--out")
    a = ap.parse_args()
    if a.worker:
        return worker(a)
    tmp = Path(tempfile.mkdtemp(prefix="b200_vbench_"))
    (tmp / "config.json").write_text(json.dumps(LLAMA_8B))
    (tmp / "generation_config.json").write_text(json.dumps({"_from_model_config": True}))
    res = {}
    for i, arm in enumerate(a.arms.split(",")):
        out = tmp / f"arm{i}.json"
        cmd = [sys.executable, __file__, "--worker", "--arm", arm, "--dir", str(tmp), "--out", str(out), "--batch", str(a.batch),
               "--prompt-len", str(a.prompt_len), "--decode", str(a.decode)]
        if a.profile:
            Path(a.profile).mkdir(parents=True, exist_ok=True)
            cmd += ["--profile", str(Path(a.profile) / f"vllm_kernels_{arm.replace(':', '_')}.txt")]
        env = dict(os.environ, PYTHONPATH=f"{ROOT}:{os.environ.get('PYTHONPATH', '')}")
        try:
            r = subprocess.run(cmd, env=env, timeout=a.timeout, capture_output=True, text=True)
        except subprocess.TimeoutExpired:
            res[arm] = {"error": f"timed out after {a.timeout} s"}
            continue
        if a.log_dir:
            Path(a.log_dir).mkdir(parents=True, exist_ok=True)
            (Path(a.log_dir) / f"vllm_log_{arm.replace(':', '_')}.txt").write_text(r.stdout + r.stderr)
        if r.returncode != 0 or not out.exists():
            res[arm] = {"error": f"rc={r.returncode}", "tail": (r.stdout + r.stderr)[-1500:]}
            continue
        res[arm] = json.loads(out.read_text())
        res[arm]["selected"] = [l.split("] ", 1)[-1][:170] for l in (r.stdout + r.stderr).splitlines()
                                if re.search(r"attention backend|Selected \S+ for|Using \S+ (backend|kernel)", l)][:6]
    ok = [k for k in res if "decode_tokens_per_s" in res[k]]
    line = {"what": "installed vLLM, Llama-3.1-8B dummy FP8 W8A8 + fp8 KV, greedy decode, CUDA graphs",
            "batch": a.batch, "prompt_len": a.prompt_len, "decode_tokens": a.decode,
            "context": f"grows {a.prompt_len} -> {a.prompt_len + a.decode} (mean ~{a.prompt_len + a.decode // 2}); NOT the fixed "
                       "4K context of bench.py",
            "arms": res}
    if "stock" in ok and "b200" in ok:
        line["b200_over_stock"] = res["b200"]["decode_tokens_per_s"] / res["stock"]["decode_tokens_per_s"]
    print(json.dumps(line))


if __name__ == "__main__":
    main()
