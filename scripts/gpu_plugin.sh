#!/bin/bash
# One gpurun call: plugin-level GPU tests + end-to-end vLLM dummy-weight greedy parity.
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
export PYTHONUNBUFFERED=1 PYTHONPATH=$PWD
timeout 900 python -m pytest tests/test_gpu_vllm_plugin.py -q -m gpu --tb=short > gpurun_out/t_plugin.log 2>&1; echo "plugin tests rc=$?"; tail -25 gpurun_out/t_plugin.log
timeout 1500 python scripts/vllm_e2e.py --model opt-125m > gpurun_out/e2e_opt.log 2>&1; echo "e2e opt rc=$?"; tail -c 3500 gpurun_out/e2e_opt.log
timeout 1500 python scripts/vllm_e2e.py --model llama-tiny --fp8 > gpurun_out/e2e_llama_fp8.log 2>&1; echo "e2e llama rc=$?"; tail -c 3500 gpurun_out/e2e_llama_fp8.log
exit 0
