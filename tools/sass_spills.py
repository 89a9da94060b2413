#!/usr/bin/env python
"""Local-memory spill instructions (STL / LDL) of one kernel by CUDA source line, from an object file built with -lineinfo.

    python tools/sass_spills.py <file.o | lib.so> <kernel-substring> [first_line last_line]
"""
import collections
import os
import re
import subprocess
import sys
import tempfile


def main():
    obj, kernel = sys.argv[1:3]
    lo, hi = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (0, 1 << 30)
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, check=True, capture_output=True)
        sass = ""
        for f in os.listdir(d):
            if f.endswith(".cubin"):
                sass += subprocess.run(["nvdisasm", "--print-line-info", os.path.join(d, f)], capture_output=True, text=True).stdout
    cur, active, cnt, tot = ("?", 0), False, collections.Counter(), 0
    for ln in sass.splitlines():
        if ln.startswith(".text."):
            active = kernel in ln
            continue
        if not active:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            if "inlined at" not in ln:
                cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.search(r"/\*[0-9a-f]{4,}\*/\s+\S", ln):
            tot += 1
            m2 = re.search(r"\b(STL|LDL)\b", ln)
            if m2 and (cur[0] != "fused.cu" or lo <= cur[1] <= hi):
                cnt[(cur, m2.group(1))] += 1
    print(f"{tot} SASS instructions, {sum(cnt.values())} spill instructions in range")
    for (k, kind), v in sorted(cnt.items()):
        print(f"  {k[0]}:{k[1]:<5d} {kind} x{v}")


if __name__ == "__main__":
    main()
