export PYTHONPATH=$PWD; mkdir -p gpurun_out
TESTS="nvfp4gemm" TAILN=3 bash scripts/gpu_iter.sh
timeout 900 python scripts/bench_kernels.py --what nvfp4 --ref 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    if 'layer' in d and d['M'] in (256,1024): print('  ', d['M'], d['layer'], d['us'], d.get('ref_us'))
"
