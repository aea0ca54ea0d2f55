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
    if args.plugin:
        from vllm_b200 import plugin

        plugin.register()
    else:
        os.environ["B200_VLLM_LINEAR"] = "0"
        os.environ["VLLM_DISABLED_KERNELS"] = "B200FP8ScaledMMLinearKernel,B200NvFp4LinearKernel"
    import torch
    from vllm import LLM, SamplingParams

    d = Path(args.dir)
    kw = dict(model=str(d), load_format="dummy", skip_tokenizer_init=True, dtype="bfloat16", max_model_len=args.max_len,
              gpu_memory_utilization=0.3, seed=0, enforce_eager=args.eager, block_size=16,
              max_num_seqs=args.num_prompts, enable_prefix_caching=False)
    if args.plugin:
        kw["attention_backend"] = "CUSTOM"
    if args.fp8:
        kw["quantization"] = "fp8"
        kw["kv_cache_dtype"] = "fp8"
    llm = LLM(**kw)
    g = torch.Generator().manual_seed(0)
    vocab = CONFIGS[args.model]["vocab_size"]
    prompts = [{"prompt_token_ids": torch.randint(10, vocab - 10, (args.prompt_len,), generator=g).tolist()}
               for _ in range(args.num_prompts)]
    sp = SamplingParams(temperature=0.0, max_tokens=args.decode, logprobs=2, ignore_eos=True, detokenize=False)
    outs = llm.generate(prompts, sp, use_tqdm=False)
    res = []
    for o in outs:
        c = o.outputs[0]
        toks = list(c.token_ids)
        # per step: logprob of the sampled token, and the top-2 {token: logprob} (to judge near-ties at a divergence)
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


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="opt-125m", choices=list(CONFIGS))
    ap.add_argument("--fp8", action="store_true")
    ap.add_argument("--eager", action="store_true")
    ap.add_argument("--num-prompts", type=int, default=32)
    ap.add_argument("--prompt-len", type=int, default=512)
    ap.add_argument("--decode", type=int, default=128)
    ap.add_argument("--max-len", type=int, default=1024)
    ap.add_argument("--timeout", type=int, default=900)
    ap.add_argument("--flat-init", action="store_true", help="keep vLLM's +-1e-3 dummy weights (flat logits)")
    ap.add_argument("--worker", action="store_true")
    ap.add_argument("--plugin", action="store_true")
    ap.add_argument("--dir")
    ap.add_argument("--out")
    args = ap.parse_args()
    if args.worker:
        return worker(args)
    tmp = Path(tempfile.mkdtemp(prefix="b200_e2e_"))
    (tmp / "config.json").write_text(json.dumps(CONFIGS[args.model]))
    (tmp / "generation_config.json").write_text(json.dumps({"_from_model_config": True}))
    runs = {}
    for name, plug in (("stock", False), ("b200", True)):
        out = tmp / f"{name}.json"
        cmd = [sys.executable, __file__, "--worker", "--model", args.model, "--dir", str(tmp), "--out", str(out),
               "--num-prompts", str(args.num_prompts), "--prompt-len", str(args.prompt_len), "--decode", str(args.decode),
               "--max-len", str(args.max_len)] + (["--plugin"] if plug else []) + (["--fp8"] if args.fp8 else []) + (
                   ["--eager"] if args.eager else []) + (["--flat-init"] if args.flat_init else [])
        env = dict(os.environ, PYTHONPATH=f"{ROOT}:{os.environ.get('PYTHONPATH', '')}")
        r = subprocess.run(cmd, env=env, timeout=args.timeout, capture_output=True, text=True)
        if r.returncode != 0 or not out.exists():
            print(json.dumps({"ok": False, "stage": name, "rc": r.returncode, "tail": (r.stdout + r.stderr)[-3000:]}))
            sys.exit(2 if name == "b200" else 3)  # 3 = the stock engine itself does not run here
        runs[name] = json.loads(out.read_text())
    a, b = runs["stock"]["results"], runs["b200"]["results"]
    n_tok = sum(len(x["tokens"]) for x in a)
    same = sum(int(t1 == t2) for x, y in zip(a, b) for t1, t2 in zip(x["tokens"], y["tokens"]))
    first_div = [next((i for i, (t1, t2) in enumerate(zip(x["tokens"], y["tokens"])) if t1 != t2), len(x["tokens"]))
                 for x, y in zip(a, b)]
    # logprob agreement on the common prefix (before any divergence the two engines saw identical contexts)
    dl = [abs(l1 - l2) for x, y, fd in zip(a, b, first_div) for l1, l2 in zip(x["logprobs"][:fd], y["logprobs"][:fd])]
    # at a stream's first divergent step both engines still saw the same context: the step is a legitimate
    # near-tie iff each engine ranks the other's choice second and within `gap` of its own
    gaps, not_in_top2 = [], 0
    for x, y, fd in zip(a, b, first_div):
        if fd == len(x["tokens"]):
            continue
        ta, tb = x["tokens"][fd], y["tokens"][fd]
        for top, mine, other in ((x["top"][fd], ta, tb), (y["top"][fd], tb, ta)):
            top = {int(k): v for k, v in top.items()}
            if other in top and mine in top:
                gaps.append(abs(top[mine] - top[other]))
            else:
                not_in_top2 += 1
    spread = [abs(l) for x in a for l in x["logprobs"]]
    rep = {"ok": True, "model": args.model, "fp8": args.fp8, "eager": args.eager, "flat_init": args.flat_init,
           "tokens": n_tok, "tokens_equal": same,
           "streams_identical": sum(int(fd == len(x["tokens"])) for x, fd in zip(a, first_div)), "streams": len(a),
           "common_prefix_tokens": len(dl),
           "max_abs_logprob_diff_common_prefix": max(dl) if dl else None,
           "mean_abs_logprob_diff_common_prefix": sum(dl) / len(dl) if dl else None,
           "divergences": sum(int(fd < len(x["tokens"])) for x, fd in zip(a, first_div)),
           "divergence_max_top2_gap": max(gaps) if gaps else None, "divergence_other_not_in_top2": not_in_top2,
           "mean_abs_logprob_stock": sum(spread) / len(spread), "flat_logprob_would_be": math.log(CONFIGS[args.model]["vocab_size"]),
           "b200": runs["b200"]["sel"], "stock": runs["stock"]["sel"]}
    print(json.dumps(rep))


if __name__ == "__main__":
    main()
