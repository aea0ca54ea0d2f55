#!/usr/bin/env python
"""SASS mnemonic counts per kernel of vllm_b200/libb200vllm.so (cuobjdump -sass): proof that the Blackwell paths are
in the binary.   python scripts/sass_counts.py > profiles/rNN_sass_counts.txt   (runs without a GPU)"""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "vllm_b200" / "libb200vllm.so"
WHAT = [("UTCQMMA", "tcgen05.mma f8f6f4 (UTCQMMA)"), ("UTCOMMA", "tcgen05.mma block-scaled mxf4nvf4 (UTCOMMA)"),
        ("UTCHMMA", "tcgen05.mma f16 (UTCHMMA)"), ("UTCCP", "tcgen05.cp (UTCCP)"), ("LDTM", "tcgen05.ld (LDTM)"),
        ("UTMALDG", "TMA load (UTMALDG)"), ("UTMASTG", "TMA store (UTMASTG)"), ("LDGMC", "multimem.ld_reduce (LDGMC)"),
        ("HMMA", "mma.sync (HMMA)"), ("STG.E.ENL2.256", "256-bit store")]


def main() -> None:
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True,
                           text=True).stdout.splitlines()
    blocks = re.split(r"\s*Function : \S+", sass)[1:]
    print("# SASS mnemonic counts per kernel of vllm_b200/libb200vllm.so (cuobjdump -sass, sm_100a), kernels with at least one")
    print("# of: " + ", ".join(k for k, _ in WHAT))
    print("# (multimem.ld_reduce...acc::f32 = LDGMC.E.HPADD.*; multimem.st has no mnemonic of its own: it is STG.E.128.STRONG.SYS")
    print("#  to the multicast address, the NVSwitch replicates it)")
    total = collections.Counter()
    for name, body in zip(names, blocks):
        short = re.sub(r"^void ", "", name)
        short = re.sub(r"\((CUtensorMap_st|b200::|unsigned|float|int|long|const|__nv|__half|void|char).*$", "", short)
        c = collections.Counter()
        for line in body.splitlines():
            m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
            if not m:
                continue
            op = m.group(1)
            for key, _ in WHAT:
                if op.startswith(key) or (key == "STG.E.ENL2.256" and op.startswith("STG") and ".256" in op):
                    c[key] += 1
                    break
        if c:
            total.update(c)
            print(f"\n{short}\n    " + ", ".join(f"{lbl}: {c[k]}" for k, lbl in WHAT if c[k]))
    print("\n# totals: " + ", ".join(f"{k} {v}" for k, v in total.items()))


if __name__ == "__main__":
    sys.exit(main())
