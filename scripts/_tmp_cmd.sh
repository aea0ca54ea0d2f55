export PYTHONPATH=$PWD; mkdir -p gpurun_out
TESTS="attn" TAILN=4 bash scripts/gpu_iter.sh
timeout 600 python scripts/bench_kernels.py --what attn > gpurun_out/kb_attn.log 2>&1; grep paged gpurun_out/kb_attn.log | cut -c1-140
