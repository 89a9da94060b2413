#!/usr/bin/env python
"""SASS opcode histogram per kernel of libngp_b200.so (evidence for what the shipped binary runs on: tcgen05 = UTCHMMA / LDTM / STTM /
UTCBAR, mma.sync = HMMA, TMA = UTMALDG / UTMASTG):   python tools/sass_opcodes.py [lib.so] > profiles/r02_sass_opcodes.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INTEREST = ("UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTCATOMSWS", "HMMA", "UTMALDG", "UTMASTG", "SYNCS", "ELECT", "LDG", "STG", "LDS", "STS", "ATOMS", "ATOMG",
            "RED", "BAR", "NANOSLEEP", "LDL", "STL", "F2FP", "HFMA2", "HMUL2", "FFMA2", "MUFU")


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "nerfsafetyvalidation_b200", "_lib", "libngp_b200.so")
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", ln)
        if m and cur:
            kernels[cur][m.group(1)] += 1
    print(f"# {os.path.relpath(lib, ROOT)}: static SASS instruction counts per kernel (cuobjdump -sass); opcodes of interest, then the total")
    for k, c in kernels.items():
        name = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip() or k
        name = re.sub(r"\(.*", "", name)
        picks = "  ".join(f"{op}={c[op]}" for op in INTEREST if c[op])
        print(f"{name:55s} total={sum(c.values()):5d}  {picks}")


if __name__ == "__main__":
    main()
