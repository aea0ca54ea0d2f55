#!/usr/bin/env python
"""Turn ncu outputs (gpurun_out/) into the small text summaries kept under profiles/.

  python scripts/ncu_summary.py launches gpurun_out/launches.csv        > profiles/rNN_launches.txt
  python scripts/ncu_summary.py full gpurun_out/prof_decode.ncu-rep     > profiles/rNN_decode_full.txt
  python scripts/ncu_summary.py traffic gpurun_out/prof_decode.ncu-rep paged_decode_kernel   > profiles/attn_traffic.json
      (per-launch dram bytes of the kernel + the digest of the kernel sources it was built from; bench.py reports
       `roofline.traffic` from that file only while the digest still matches the tree)
"""
import collections
import csv
import re
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_gen5_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__waves_per_multiprocessor", "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__warps_eligible.avg.per_cycle_active",
]
STALL = "smsp__average_warps_issue_stalled_"


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg, tot = collections.OrderedDict(), 0.0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(row["Metric Unit"], 1.0)
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        name = re.sub(r"^void ", "", name)[:100]
        key = (name, row.get("Grid Size", ""))
        a = agg.setdefault(key, [0, 0.0])
        a[0] += 1
        a[1] += v
        tot += v
    print(f"# {path}: per-launch device time under ncu (cold cache, serialised): compare SHARES\n# total {tot:.1f} us")
    print(f"{'total_us':>10} {'n':>4} {'avg_us':>9} {'share':>6}  kernel  grid")
    for (k, grid), (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{t:10.1f} {n:4d} {t / n:9.2f} {100 * t / tot:5.1f}%  {k}  {grid}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("kernel:", r[hdr.index("Kernel Name")][:140])
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"  {k:80s} {r[i]:>16s} {units[i]}")
        st = [(float(r[i] or 0), k[len(STALL):-len('_per_issue_active.ratio')]) for i, k in enumerate(hdr)
              if k.startswith(STALL) and k.endswith("_per_issue_active.ratio")]
        print("  stalls (warps per issue-active cycle):", ", ".join(f"{n}={v:.2f}" for v, n in sorted(st, reverse=True)[:8]))
        print()


def kernel_src_digest() -> str:
    """sha256 over the attention kernel sources (same function in bench.py)."""
    import hashlib
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent / "vllm_b200" / "csrc"
    h = hashlib.sha256()
    for f in ("paged_decode.cu", "paged_decode.cuh", "common.cuh"):
        h.update((root / f).read_bytes())
    return h.hexdigest()[:16]


def traffic(path, pattern):
    import json

    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    vals = []
    for r in rows[2:]:
        if pattern not in r[hdr.index("Kernel Name")]:
            continue
        tot = 0.0
        for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = hdr.index(k)
            tot += float(r[i].replace(",", "")) * scale.get(units[i], 1.0)
        vals.append(tot)
    print(json.dumps({"kernel": pattern, "launches": len(vals), "dram_bytes_per_launch": sum(vals) / max(len(vals), 1),
                      "per_launch": vals, "source": path, "kernel_src_sha16": kernel_src_digest(),
                      "workload": "configs[2] TP1: batch 128 x ctx 4096, fp8 KV, 32q/8kv, D=128"}))


if __name__ == "__main__":
    if sys.argv[1] == "traffic":
        traffic(sys.argv[2], sys.argv[3])
    else:
        {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
