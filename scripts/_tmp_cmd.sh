export PYTHONPATH=$PWD; mkdir -p gpurun_out
TESTS="attn quant" TAILN=3 bash scripts/gpu_iter.sh
python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | cut -c1-330
