#!/usr/bin/env python
"""stdin: `ncu --page source --csv --print-source sass`; stdout: per kernel the N most-sampled SASS
instructions (where the warps stall)."""
import csv
import sys

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10
rows = list(csv.reader(sys.stdin))
kernel, hdr, data = None, None, []
seen = set()


def flush():
    if kernel is None or hdr is None or not data or kernel in seen:
        return
    seen.add(kernel)
    i_s, i_src = hdr.index("# Samples"), hdr.index("Source")
    tot = sum(int(r[i_s]) for r in data) or 1
    print(f"== {kernel[:110]}  ({tot} samples, {len(data)} instructions)")
    for r in sorted(data, key=lambda r: -int(r[i_s]))[:n]:
        print(f"   {100 * int(r[i_s]) / tot:5.1f}%  {r[i_src].strip()[:100]}")


for r in rows:
    if r and r[0] == "Kernel Name":
        flush()
        kernel, hdr, data = r[1], None, []
    elif r and r[0] == "Address":
        hdr = r
    elif hdr and len(r) == len(hdr):
        data.append(r)
flush()
