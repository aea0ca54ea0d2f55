#!/bin/bash
# One gpurun call: smoke, GPU parity tests (each group in its own process with a timeout so a
# faulting kernel cannot take the rest down), then a short bench.  Logs -> gpurun_out/.
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
export PYTHONUNBUFFERED=1
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1

run() { # name timeout cmd...
  local name=$1 to=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout "$to" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?
  echo "rc=$rc" | tee -a gpurun_out/summary.txt
  tail -n 15 "gpurun_out/$name.log" | tee -a gpurun_out/summary.txt
  return $rc
}

: > gpurun_out/summary.txt
run smoke 600 python __graft_entry__.py smoke
run t_cache 600 python -m pytest tests/test_gpu_attention.py -q -m gpu -k "cache_write or slot_mapping or merge" -x --tb=short
run t_decode 900 python -m pytest tests/test_gpu_attention.py -q -m gpu -k "decode" --tb=short
run t_fp8quant 600 python -m pytest tests/test_gpu_quant_gemm.py -q -m gpu -k "fp8_quant or elementwise" --tb=short
run t_nvfp4quant 600 python -m pytest tests/test_gpu_quant_gemm.py -q -m gpu -k "nvfp4_quant" --tb=short
run t_fp8gemm 600 python -m pytest tests/test_gpu_quant_gemm.py -q -m gpu -k "fp8_gemm or fp8_linear" --tb=short
run t_nvfp4gemm 600 python -m pytest tests/test_gpu_quant_gemm.py -q -m gpu -k "nvfp4_gemm" --tb=short
if [ "${SKIP_BENCH:-0}" != "1" ]; then
  run bench 900 python bench.py --steps 10 --warmup 3
fi
echo "=== done" >> gpurun_out/summary.txt
exit 0
