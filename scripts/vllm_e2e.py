#!/usr/bin/env python
"""End-to-end drop-in check inside an UNMODIFIED vLLM (the installed wheel), GPU box only.

Runs `vllm.LLM` on dummy weights (`load_format="dummy"`, deterministic per-parameter seed 1234,
vllm/model_executor/model_loader/weight_utils.py:1289-1363) with synthetic prompt token ids, once with
vLLM's stock attention backend / linear kernels and once with the B200 plugin selected
(`attention_backend="CUSTOM"`, B200 linear kernels first), greedy decoding, and compares the token
streams and per-step logprobs.  One engine per subprocess so GPU memory is returned in between.

  python scripts/vllm_e2e.py --model opt-125m            # BASELINE configs[0] (shape of facebook/opt-125m)
  python scripts/vllm_e2e.py --model llama-tiny --fp8    # Llama-style GQA + online FP8 W8A8 + fp8 KV cache

Model configs are written locally (no network): the architecture constants are the public ones.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import re
import subprocess
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent

CONFIGS = {
    "opt-125m": {
        "architectures": ["OPTForCausalLM"], "model_type": "opt", "hidden_size": 768, "num_hidden_layers": 12,
        "num_attention_heads": 12, "ffn_dim": 3072, "vocab_size": 50272, "max_position_embeddings": 2048,
        "word_embed_proj_dim": 768, "do_layer_norm_before": True, "activation_function": "relu",
        "pad_token_id": 1, "bos_token_id": 2, "eos_token_id": 2, "torch_dtype": "bfloat16", "layer_norm_elementwise_affine": True,
        "enable_bias": True, "dropout": 0.0, "attention_dropout": 0.0, "layerdrop": 0.0, "init_std": 0.02, "prefix": "</s>",
        "_remove_final_layer_norm": False, "use_cache": True,
    },
    "llama-tiny": {  # Llama-3.1 architecture, GQA 4:1, D=128, scaled down so the check runs in a minute
        "architectures": ["LlamaForCausalLM"], "model_type": "llama", "hidden_size": 1024, "intermediate_size": 3584,
        "num_hidden_layers": 4, "num_attention_heads": 8, "num_key_value_heads": 2, "head_dim": 128, "vocab_size": 32000,
        "max_position_embeddings": 8192, "rms_norm_eps": 1e-5, "rope_theta": 500000.0, "hidden_act": "silu",
        "tie_word_embeddings": False, "torch_dtype": "bfloat16", "bos_token_id": 1, "eos_token_id": 2,
        "attention_bias": False, "mlp_bias": False,
    },
}


def patch_dummy_init() -> None:
    """vLLM's dummy loader fills every float parameter with uniform(-1e-3, 1e-3) from the SAME seed
    (weight_utils.py initialize_single_dummy_weight): logits come out flat, every logprob is -log(vocab) and a
    comparison of two engines is insensitive.  Both arms therefore run with this replacement:
    matrices ~ uniform(+-sqrt(3 / fan_in)) (variance-preserving), vectors (norm weights, biases) ~ uniform(0.5, 1.5),
    one seed per parameter in traversal order - deterministic, identical in both arms, logits with unit spread."""
    import torch
    from vllm.model_executor.model_loader import dummy_loader, weight_utils

    counter = [0]

    @torch.no_grad()
    def init(param, low=-1e-3, high=1e-3, seed=1234):
        if param.device.type == "meta" or not torch.is_floating_point(param):
            return
        counter[0] += 1
        g = torch.Generator(device=param.data.device)
        g.manual_seed(seed + counter[0])
        if param.dim() >= 2:
            a = (3.0 / param.shape[-1]) ** 0.5
            lo, hi = -a, a
        else:
            lo, hi = 0.5, 1.5
        if torch.finfo(param.data.dtype).bits < 16:
            param.data.copy_(param.data.to(torch.float16).uniform_(lo, hi, generator=g).to(param.data.dtype))
        else:
            param.uniform_(lo, hi, generator=g)

    weight_utils.initialize_single_dummy_weight = init
    dummy_loader.initialize_single_dummy_weight = init


def worker(args) -> None:
    sys.path.insert(0, str(ROOT))
    os.environ.setdefault("VLLM_ENABLE_V1_MULTIPROCESSING", "0")
    if not args.flat_init:
        patch_dummy_init()
    # arms: stock[:BACKEND] | b200 (our attention + our linears) | b200lin:BACKEND (our linears, a stock attention
    #       backend) | b200attn (our attention, stock linears)
    kind = args.arm.split(":", 1)[0]
    plugin_on = kind.startswith("b200")
    backend = args.arm.split(":", 1)[1] if ":" in args.arm else None  # stock:TRITON_ATTN, stock:FLASHINFER, ...
    if plugin_on:
        if kind == "b200attn":
            os.environ["B200_VLLM_LINEAR"] = "0"
            os.environ["VLLM_DISABLED_KERNELS"] = "B200FP8ScaledMMLinearKernel,B200NvFp4LinearKernel"
        from vllm_b200 import plugin

        plugin.register()
        if kind != "b200lin":
            backend = "CUSTOM"
    else:
        os.environ["B200_VLLM_LINEAR"] = "0"
        os.environ["VLLM_DISABLED_KERNELS"] = "B200FP8ScaledMMLinearKernel,B200NvFp4LinearKernel"
    import torch
    from vllm import LLM, SamplingParams

    d = Path(args.dir)
    kw = dict(model=str(d), load_format="dummy", skip_tokenizer_init=True, dtype="bfloat16", max_model_len=args.max_len,
              gpu_memory_utilization=0.3, seed=0, enforce_eager=args.eager, block_size=16,
              max_num_seqs=args.num_prompts, enable_prefix_caching=False)
    if backend:
        kw["attention_backend"] = backend
    if args.fp8 or args.w_fp8:
        kw["quantization"] = "fp8"
    if args.fp8 or args.kv_fp8:
        kw["kv_cache_dtype"] = "fp8"
    llm = LLM(**kw)
    g = torch.Generator().manual_seed(0)
    vocab = CONFIGS[args.model]["vocab_size"]
    prompts = [{"prompt_token_ids": torch.randint(10, vocab - 10, (args.prompt_len,), generator=g).tolist()}
               for _ in range(args.num_prompts)]
    sp = SamplingParams(temperature=0.0, max_tokens=args.decode, logprobs=5, ignore_eos=True, detokenize=False)
    outs = llm.generate(prompts, sp, use_tqdm=False)
    res = []
    for o in outs:
        c = o.outputs[0]
        toks = list(c.token_ids)
        # per step: logprob of the sampled token, and the top-5 {token: logprob} (to judge near-ties at a divergence)
        lps = [float(lp[t].logprob) for lp, t in zip(c.logprobs, toks)]
        top = [{int(k): float(v.logprob) for k, v in lp.items()} for lp in c.logprobs]
        res.append({"tokens": toks, "logprobs": lps, "top": top})
    sel = {}
    try:
        import vllm_b200._lib as L

        sel["b200_lib_loaded"] = L.is_loaded()
        sel["b200_launches"] = L.num_kernels_launched() if L.is_loaded() else 0
    except Exception:
        sel["b200_lib_loaded"] = False
    Path(args.out).write_text(json.dumps({"results": res, "sel": sel}))


def compare(a, b, gap_tol: float) -> dict:
    """Two arms' greedy streams: logprob agreement on each stream's common prefix (identical contexts) and, at the
    first divergent step (contexts still identical), how close a call it was in EACH engine's own top-5."""
    first_div = [next((i for i, (t1, t2) in enumerate(zip(x["tokens"], y["tokens"])) if t1 != t2), len(x["tokens"]))
                 for x, y in zip(a, b)]
    dl = [abs(l1 - l2) for x, y, fd in zip(a, b, first_div) for l1, l2 in zip(x["logprobs"][:fd], y["logprobs"][:fd])]
    gaps, missing = [], 0
    for x, y, fd in zip(a, b, first_div):
        if fd == len(x["tokens"]):
            continue
        ta, tb = x["tokens"][fd], y["tokens"][fd]
        for top, mine, other in ((x["top"][fd], ta, tb), (y["top"][fd], tb, ta)):
            top = {int(k): v for k, v in top.items()}
            if other in top and mine in top:
                gaps.append(abs(top[mine] - top[other]))
            else:
                missing += 1
    dl_sorted = sorted(dl)
    return {"tokens": sum(len(x["tokens"]) for x in a),
            "tokens_equal": sum(int(t1 == t2) for x, y in zip(a, b) for t1, t2 in zip(x["tokens"], y["tokens"])),
            "streams": len(a), "streams_identical": sum(int(fd == len(x["tokens"])) for x, fd in zip(a, first_div)),
            "common_prefix_tokens": len(dl),
            "max_abs_logprob_diff": max(dl) if dl else None, "mean_abs_logprob_diff": sum(dl) / len(dl) if dl else None,
            "p90_abs_logprob_diff": dl_sorted[int(0.9 * (len(dl) - 1))] if dl else None,
            "divergences": sum(int(fd < len(x["tokens"])) for x, fd in zip(a, first_div)),
            "divergence_max_top_gap": max(gaps) if gaps else None,
            "divergence_mean_top_gap": sum(gaps) / len(gaps) if gaps else None,
            "divergence_other_not_in_top5": missing,
            "divergences_beyond_gap_tol": sum(int(g > gap_tol) for g in gaps)}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="opt-125m", choices=list(CONFIGS))
    ap.add_argument("--fp8", action="store_true", help="FP8 W8A8 linears AND fp8 KV cache")
    ap.add_argument("--w-fp8", action="store_true", help="FP8 W8A8 linears only")
    ap.add_argument("--kv-fp8", action="store_true", help="fp8 KV cache only")
    ap.add_argument("--eager", action="store_true")
    ap.add_argument("--num-prompts", type=int, default=32)
    ap.add_argument("--prompt-len", type=int, default=512)
    ap.add_argument("--decode", type=int, default=128)
    ap.add_argument("--max-len", type=int, default=1024)
    ap.add_argument("--timeout", type=int, default=900)
    ap.add_argument("--flat-init", action="store_true", help="keep vLLM's +-1e-3 dummy weights (flat logits)")
    ap.add_argument("--arms", default="stock,b200",
                    help="comma list of engines to run: stock | stock:<ATTENTION_BACKEND> | b200; every arm is compared "
                         "with the first one")
    ap.add_argument("--gap-tol", type=float, default=0.5)
    ap.add_argument("--worker", action="store_true")
    ap.add_argument("--arm", default="stock")
    ap.add_argument("--dir")
    ap.add_argument("--out")
    args = ap.parse_args()
    if args.worker:
        return worker(args)
    tmp = Path(tempfile.mkdtemp(prefix="b200_e2e_"))
    (tmp / "config.json").write_text(json.dumps(CONFIGS[args.model]))
    (tmp / "generation_config.json").write_text(json.dumps({"_from_model_config": True}))
    arms = args.arms.split(",")
    runs, failed = {}, {}
    for i, arm in enumerate(arms):
        out = tmp / f"arm{i}.json"
        cmd = [sys.executable, __file__, "--worker", "--arm", arm, "--model", args.model, "--dir", str(tmp), "--out", str(out),
               "--num-prompts", str(args.num_prompts), "--prompt-len", str(args.prompt_len), "--decode", str(args.decode),
               "--max-len", str(args.max_len)]
        for flag in ("fp8", "w_fp8", "kv_fp8", "eager", "flat_init"):
            if getattr(args, flag):
                cmd.append("--" + flag.replace("_", "-"))
        env = dict(os.environ, PYTHONPATH=f"{ROOT}:{os.environ.get('PYTHONPATH', '')}")
        r = subprocess.run(cmd, env=env, timeout=args.timeout, capture_output=True, text=True)
        if r.returncode != 0 or not out.exists():
            failed[arm] = {"rc": r.returncode, "tail": (r.stdout + r.stderr)[-2500:]}
            if i == 0 or arm.startswith("b200"):
                print(json.dumps({"ok": False, "stage": arm, **failed[arm]}))
                sys.exit(2 if arm.startswith("b200") else 3)  # 3 = the stock engine itself does not run here
            continue
        runs[arm] = json.loads(out.read_text())
        # which kernels the engine picked, from its own log lines
        runs[arm]["sel"]["log"] = [l.split("] ", 1)[-1][:160] for l in (r.stdout + r.stderr).splitlines()
                                   if re.search(r"attention backend|Selected \S+ for|Using \S+ (backend|kernel)", l)][:8]
    base = runs[arms[0]]["results"]
    spread = [abs(l) for x in base for l in x["logprobs"]]
    rep = {"ok": True, "model": args.model, "w_fp8": args.fp8 or args.w_fp8, "kv_fp8": args.fp8 or args.kv_fp8,
           "eager": args.eager, "flat_init": args.flat_init, "base": arms[0],
           "mean_abs_logprob_base": sum(spread) / len(spread), "flat_logprob_would_be": math.log(CONFIGS[args.model]["vocab_size"]),
           "pairs": {arm: compare(base, runs[arm]["results"], args.gap_tol) for arm in arms[1:] if arm in runs},
           "sel": {arm: runs[arm]["sel"] for arm in runs}, "arms_failed": failed}
    print(json.dumps(rep))


if __name__ == "__main__":
    main()
