"""north_star condition (3): greedy-decode token streams and per-step logprobs of an UNMODIFIED installed vLLM with
the B200 plugin selected (attention backend CUSTOM + B200 linear kernels) against the same engine with its stock
kernels, on identical dummy-weight models (scripts/vllm_e2e.py runs the engines, one per subprocess).

The dummy weights are re-initialised with unit-spread values in EVERY arm (vLLM's own +-1e-3 fill gives flat
logits - every logprob equals -log(vocab) and the comparison would be blind).  With informative
logits two correct engines part ways at genuine near-ties, and with FP8 activations (a discontinuous per-token
quantiser in front of every linear) small differences are amplified: measured on B200, the reference's OWN two
attention backends (FlashInfer trtllm-gen vs Triton, identical CUTLASS linears) differ by 0.10 mean / 0.28 max in the
sampled token's logprob and diverge on all 16 streams within a few steps.  The bars therefore are:

  * bf16 model (no quantiser): absolute - logprobs on the common prefix within LP_TOL_BF16, every divergence a
    near-tie inside both engines' top-5;
  * fp8 model: RELATIVE to the reference's own spread measured in the same run (stock vs stock:TRITON_ATTN):
    ours-vs-stock may not exceed SPREAD_FACTOR x that spread (+ a small floor), plus absolute caps;
  * isolation: our FP8 linears under the STOCK Triton attention vs stock linears under the same attention - only
    the GEMM differs, and the distance is STILL ~0.1 (the quantiser amplifies one-ulp differences): this measures
    the FP8 model's noise floor and carries absolute caps only;
  * in every case the plugin really ran (its library launched kernels, vLLM's log names our classes) and the
    logits were not flat.
Reference for the method: tests/utils.py:1180 compare_two_settings, tests/v1/attention/test_attention_backends.py.
"""
import json
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.gpu

LP_TOL_BF16 = 0.08    # bf16 logits: one ulp is 0.03-0.06 at |logit| ~ 8 (measured max 0.063, mean 0.009)
TIE_TOL_BF16 = 0.15   # top-1 vs the other engine's token at a divergence (measured max 0.031)
SPREAD_FACTOR = 1.6   # ours-vs-stock <= factor x (stock-vs-stock) + floor
SIZES = ("--num-prompts", "16", "--prompt-len", "256", "--decode", "48")


def _run(*flags: str) -> dict:
    try:
        import vllm  # noqa: F401
    except Exception as e:  # pragma: no cover
        pytest.skip(f"vllm is not importable here: {e}")
    r = subprocess.run([sys.executable, str(ROOT / "scripts" / "vllm_e2e.py"), *flags, *SIZES], capture_output=True,
                       text=True, timeout=2400)
    line = next((l for l in reversed(r.stdout.strip().splitlines()) if l.startswith("{")), None)
    assert line, f"no report (rc={r.returncode}): {(r.stdout + r.stderr)[-3000:]}"
    rep = json.loads(line)
    if r.returncode == 3:  # the STOCK engine does not run on this box: nothing to compare against
        pytest.skip(f"stock vLLM engine failed here: {rep.get('tail', '')[-400:]}")
    assert rep.get("ok"), rep
    print(json.dumps(rep))
    return rep


def _common(rep: dict, arm: str, *, attn: bool, linear: bool) -> dict:
    sel = rep["sel"][arm]
    assert sel["b200_lib_loaded"] and sel["b200_launches"] > 0, "the plugin path did not run our kernels"
    log = " ".join(sel.get("log", []))
    if attn:
        assert "CUSTOM" in log, log
    if linear:
        assert "B200FP8ScaledMMLinearKernel" in log, log
    assert not rep["sel"][rep["base"]]["b200_lib_loaded"]
    # informative logits: the sampled tokens' mean |logprob| is far from the flat value log(vocab)
    assert rep["mean_abs_logprob_base"] < rep["flat_logprob_would_be"] - 1.0, rep
    return rep["pairs"][arm]


def test_llama_bf16_greedy_parity():
    """Llama-3.1 architecture (GQA 4:1, D=128), bf16 weights + bf16 KV, CUDA graphs on: no quantiser anywhere, so
    the comparison is absolute."""
    rep = _run("--model", "llama-tiny", "--arms", "stock,b200")
    p = _common(rep, "b200", attn=True, linear=False)
    assert p["common_prefix_tokens"] >= p["tokens"] // 4, p
    assert p["max_abs_logprob_diff"] <= LP_TOL_BF16, p
    assert p["divergence_other_not_in_top5"] == 0, p
    assert p["divergences"] == 0 or p["divergence_max_top_gap"] <= TIE_TOL_BF16, p


def test_opt125m_bf16_greedy_parity():
    """BASELINE configs[0] shape (facebook/opt-125m: MHA, D=64, bf16 KV, block 16), eager mode."""
    rep = _run("--model", "opt-125m", "--eager", "--arms", "stock,b200")
    p = _common(rep, "b200", attn=True, linear=False)
    assert p["common_prefix_tokens"] >= p["tokens"] // 4, p
    assert p["max_abs_logprob_diff"] <= LP_TOL_BF16, p
    assert p["divergence_other_not_in_top5"] == 0, p
    assert p["divergences"] == 0 or p["divergence_max_top_gap"] <= TIE_TOL_BF16, p


def test_llama_fp8_w8a8_fp8_kv_within_reference_spread():
    """configs[2]'s numerics (online FP8 W8A8 per-tensor weights + fp8 KV cache) on the Llama architecture: the
    plugin's distance from the stock engine is bounded by the stock engine's own backend-to-backend distance."""
    rep = _run("--model", "llama-tiny", "--fp8", "--arms", "stock,stock:TRITON_ATTN,b200")
    ours = _common(rep, "b200", attn=True, linear=True)
    if "stock:TRITON_ATTN" not in rep["pairs"]:
        pytest.skip(f"second stock backend did not run here: {rep['arms_failed']}")
    ref = rep["pairs"]["stock:TRITON_ATTN"]
    assert ref["common_prefix_tokens"] > 0 and ours["common_prefix_tokens"] > 0
    for key, floor in (("mean_abs_logprob_diff", 0.02), ("p90_abs_logprob_diff", 0.04), ("max_abs_logprob_diff", 0.08)):
        assert ours[key] <= SPREAD_FACTOR * ref[key] + floor, (key, ours[key], ref[key])
    # a COUNT over 16 streams that all diverge (measured: reference 3 / ours 3-6 from run to run - which near-tie a
    # stream leaves at depends on one-ulp differences): bounded relative to the reference's own count, and by half
    # of the streams
    assert ours["divergence_other_not_in_top5"] <= max(2 * ref["divergence_other_not_in_top5"] + 4, ours["streams"] // 2), (ours, ref)
    assert ours["common_prefix_tokens"] >= 0.5 * ref["common_prefix_tokens"], (ours, ref)
    # absolute caps: a wrong scale / head mapping / cache layout moves logprobs by whole units
    assert ours["mean_abs_logprob_diff"] < 0.3 and ours["max_abs_logprob_diff"] < 1.0, ours


def test_fp8_linear_kernel_in_situ_vs_cutlass():
    """Only the FP8 GEMM differs: our linear kernels under the stock Triton attention vs the stock CUTLASS kernels
    under the same attention (bf16 KV).  Measured on B200: even this - same fp8 operands, fp32 accumulation on both
    sides, so differences of one bf16 rounding in isolated outputs - ends at 0.095 mean / 0.35 max logprob
    difference, because the dynamic per-token FP8 quantiser in front of the next linear re-buckets a whole row
    when its amax moves.  So the FP8 model's noise floor IS ~0.1 whatever differs, this test can only carry
    absolute caps (a wrong scale or a transposed tile moves logprobs by whole units), and the sharp check of
    b200_scaled_mm_fp8 against CUTLASS is the kernel-level one (test_fp8_linear_kernel_against_installed_cutlass)."""
    rep = _run("--model", "llama-tiny", "--w-fp8", "--arms", "stock:TRITON_ATTN,b200lin:TRITON_ATTN")
    p = _common(rep, "b200lin:TRITON_ATTN", attn=False, linear=True)
    assert p["mean_abs_logprob_diff"] < 0.2 and p["max_abs_logprob_diff"] < 0.8, p
    assert p["divergence_other_not_in_top5"] <= 2, p
