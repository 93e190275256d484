#!/usr/bin/env python
"""SASS opcode histogram of the shipped library, per kernel family: the evidence that the TMA / mbarrier /
cluster code paths exist in the binary the tests load (cuobjdump -sass cfd_b200/libcfd_b200.so).

    python tools/sass_histogram.py > profiles/r2_sass_histogram.md
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cfd_b200", "libcfd_b200.so")
# mnemonics that prove a hardware path (B200_PROFILING.md) + the FP64 pipe
WATCH = ["UTMALDG", "UTMASTG", "UTMAPF", "SYNCS", "STAS", "UCGABAR", "CCTL", "MEMBAR", "LDGSTS", "DFMA", "DADD", "DMUL",
         "DSETP", "MUFU.RCP64H", "LDG", "STG", "LDS", "STS", "BAR", "ATOM", "RED", "NANOSLEEP", "UTCBAR", "UTCHMMA", "HMMA"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    fam, cur = collections.OrderedDict(), None
    total = collections.Counter()
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name).replace("void ", "").replace("cfdb::", "").replace("(anonymous namespace)::", "")
            cur = fam.setdefault(name, collections.Counter())
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if m and cur is not None:
            op = m.group(1)
            cur["_n"] += 1
            for w in WATCH:
                if op == w or op.startswith(w + ".") or op.startswith(w + "_"):
                    cur[w] += 1
                    total[w] += 1
    print("# SASS opcode histogram of cfd_b200/libcfd_b200.so (sm_100a), `cuobjdump -sass`, per kernel\n")
    print("Whole library: " + ", ".join("%s %d" % (w, total[w]) for w in WATCH if total[w]) + "\n")
    cols = [w for w in WATCH if total[w]]
    print("| kernel | instr | " + " | ".join(cols) + " |")
    print("|---|---|" + "---|" * len(cols))
    for name, c in fam.items():
        print("| `%s` | %d | " % (name[:70], c["_n"]) + " | ".join(str(c[w]) if c[w] else "" for w in cols) + " |")


if __name__ == "__main__":
    sys.exit(main())
