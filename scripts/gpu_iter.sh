#!/bin/bash
# One gpurun call for the tuning loop: GEMM parity tests, plugin tests, kernel sweep, short bench.
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
export PYTHONUNBUFFERED=1 PYTHONPATH=$PWD
run() { local name=$1 to=$2; shift 2; echo "=== $name"; timeout "$to" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; tail -n ${TAILN:-6} "gpurun_out/$name.log"; }
for t in ${TESTS:-fp8gemm}; do
  case $t in
    fp8gemm) run t_fp8gemm 600 python -m pytest tests/test_gpu_quant_gemm.py -q -m gpu -k "fp8_gemm or fp8_linear" --tb=short -x;;
    nvfp4gemm) run t_nvfp4gemm 600 python -m pytest tests/test_gpu_quant_gemm.py -q -m gpu -k "nvfp4_gemm" --tb=short -x;;
    quant) run t_quant 600 python -m pytest tests/test_gpu_quant_gemm.py -q -m gpu -k "quant or elementwise" --tb=short;;
    attn) run t_attn 900 python -m pytest tests/test_gpu_attention.py -q -m gpu --tb=short;;
    plugin) run t_plugin 900 python -m pytest tests/test_gpu_vllm_plugin.py -q -m gpu --tb=short;;
    e2e_llama) run e2e_llama 1500 python scripts/vllm_e2e.py --model llama-tiny --fp8;;
    e2e_opt) run e2e_opt 1500 python scripts/vllm_e2e.py --model opt-125m;;
    all) run t_all 1800 python -m pytest tests -q -m gpu --tb=short;;
  esac
done
if [ -n "${KBENCH:-}" ]; then TAILN=80 run kernels 1200 python scripts/bench_kernels.py --what "$KBENCH" ${KREF:+--ref}; fi
if [ "${BENCH:-0}" = "1" ]; then TAILN=3 run bench 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline; fi
exit 0
