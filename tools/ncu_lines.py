#!/usr/bin/env python
"""Aggregates ncu's per-SASS-instruction warp-stall samples to CUDA source lines.

    ncu -i prof.ncu-rep --page source --csv > src.csv
    cuobjdump -xelf all lib.so; nvdisasm --print-line-info X.cubin > all.sass
    python tools/ncu_lines.py src.csv all.sass <kernel-substring> [top]

ncu's CSV source page carries no line correlation, so instructions are matched by order with nvdisasm's listing
(`//## File "...", line N` annotations; inlined frames keep the innermost line)."""
import csv
import re
import sys
from collections import defaultdict


def sass_lines(path, kernel):
    out, cur, active = [], ("?", 0), False
    for ln in open(path):
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
            out.append(cur)
    return out


def main():
    src_csv, sass, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    sort_key = sys.argv[5] if len(sys.argv) > 5 else "samples"   # or "inst"
    rows = list(csv.reader(open(src_csv)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    col = {h: i for i, h in enumerate(hdr)}
    insts = [r for r in rows[hdr_i + 1:] if len(r) == len(hdr)]
    lines = sass_lines(sass, kernel)
    print(f"# {len(insts)} instructions in the ncu page, {len(lines)} in the nvdisasm listing", file=sys.stderr)
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = defaultdict(lambda: defaultdict(float))
    total = 0.0
    for k, r in enumerate(insts):
        key = lines[k] if k < len(lines) else ("?", 0)
        n = float(r[col["# Samples"]] or 0)
        total += n
        agg[key]["samples"] += n
        agg[key]["inst"] += float(r[col["Instructions Executed"]] or 0)
        agg[key]["tinst"] += float(r[col["Thread Instructions Executed"]] or 0)
        for s in stalls:
            agg[key][s] += float(r[col[s]] or 0)
    ti = sum(a["inst"] for a in agg.values())
    print(f"total samples {total:.0f}, warp instructions {ti:.4g}, thread instructions {sum(a['tinst'] for a in agg.values()):.4g}")
    for key, a in sorted(agg.items(), key=lambda kv: -kv[1][sort_key])[:top]:
        tops = sorted(((a[s], s[6:]) for s in stalls), reverse=True)[:3]
        print(f"{key[0]}:{key[1]:<5d} {100 * a['samples'] / total:5.1f}%  inst {a['inst']:.3g} ({100 * a['inst'] / ti:.1f}%, {a['tinst'] / max(a['inst'], 1):.0f} thr)  " + "  ".join(f"{n}={100 * v / max(a['samples'], 1):.0f}%" for v, n in tops))


if __name__ == "__main__":
    main()
