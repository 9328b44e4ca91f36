#!/usr/bin/env python
"""SASS opcode evidence per kernel of libhgpumap.so: how often the Blackwell-native mnemonics occur (tcgen05.mma =
UTC*MMA, tcgen05.ld = LDTM, TMA = UTMALDG / UBLKCP, tcgen05.commit = UTCBAR), the shared-memory atomics of the d = 2
combiners, FP64 reductions, 128-bit streaming loads, and the legacy tensor path (HMMA: must stay 0).
    python tools/sass_histogram.py > profiles/rNN_sass_histogram.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "hadoop-1.0.3-gpu_b200", "libhgpumap.so")
COLS = [("UTC*MMA", r"^UTC\w*MMA"), ("LDTM", r"^LDTM"), ("UTMALDG", r"^UTMALDG"), ("UBLKCP", r"^UBLKCP"),
        ("UTCBAR", r"^UTCBAR"), ("ATOMS.ADD/INC", r"^ATOMS\.(ADD|POPC)"), ("ATOMS.CAS", r"^ATOMS\.CAS"),
        ("REDG .F64", r"^(REDG|ATOMG)\..*F64"), ("LDG.128", r"^LDG\..*128"), ("LDS.U8", r"^LDS\.U8"), ("HMMA", r"^HMMA")]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    name = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            kernels[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][\w.]*)", line)
        if m and name:
            op = m.group(1)
            kernels[name]["instructions"] += 1
            for col, pat in COLS:
                if re.search(pat, op):
                    kernels[name][col] += 1
    names = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    rows = []
    for mangled, pretty in zip(kernels, names):
        pretty = re.sub(r"\(anonymous namespace\)::|hgm::|\(int\)|\(bool\)|void ", "", pretty)
        pretty = re.sub(r"\((const |unsigned |float|double|long|int|CUtensorMap|Grid2D|FxParams|Tensor|cudaTextureObject).*$", "", pretty)
        rows.append((pretty, kernels[mangled]))
    rows.sort(key=lambda r: r[0])
    print(f"# SASS opcode histogram per kernel (`{os.path.basename(LIB)}`, `cuobjdump -sass`, sm_100a)\n")
    print("Counts of static instructions.  `UTC*MMA` = tcgen05.mma, `LDTM` = tcgen05.ld, `UTMALDG` = TMA tensor load, `UBLKCP` = TMA bulk copy,")
    print("`UTCBAR` = tcgen05.commit; `HMMA` (the legacy mma.sync path) must be 0 everywhere.\n")
    print("| kernel | instructions | " + " | ".join(c for c, _ in COLS) + " |")
    print("|---|---|" + "---|" * len(COLS))
    for pretty, c in rows:
        print(f"| `{pretty}` | {c['instructions']} | " + " | ".join(str(c[col]) for col, _ in COLS) + " |")
    total = collections.Counter()
    for _, c in rows:
        total.update(c)
    print(f"| **all {len(rows)} kernels** | {total['instructions']} | " + " | ".join(str(total[col]) for col, _ in COLS) + " |")


if __name__ == "__main__":
    main()
