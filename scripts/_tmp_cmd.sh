export PYTHONPATH=$PWD; mkdir -p gpurun_out
TESTS="all" TAILN=15 bash scripts/gpu_iter.sh
for pdl in 1 0; do
echo "=== PDL=$pdl kernels"
B200_PDL=$pdl timeout 600 python scripts/bench_kernels.py --what gemm --ms 16,128,256 > gpurun_out/kb_pdl$pdl.log 2>&1
python - <<PY
import json
for l in open("gpurun_out/kb_pdl$pdl.log"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    if "layer" in d: print(d["M"], d["layer"], d["us"], d.get("ref_us"))
PY
echo "=== PDL=$pdl bench"
B200_PDL=$pdl timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | cut -c1-400
done
