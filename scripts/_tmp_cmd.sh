export PYTHONPATH=$PWD; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_quant_gemm.py tests/test_gpu_attention.py -q -m gpu -k "rotary or rope or fused or cache" --tb=short 2>&1 | tail -4
python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | cut -c1-330
