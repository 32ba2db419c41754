#!/usr/bin/env python
"""stdin/arg: an `ncu --metrics gpu__time_duration.sum --csv` log of one training step; prints per kernel name the
summed device time, share and launch count (cold-cache, serialised: compare SHARES, not absolutes)."""
import collections
import csv
import re
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 60
rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 10]
hdr = next((r for r in rows if "Kernel Name" in r), None)
if hdr is None:
    sys.exit("no ncu csv rows in " + path)
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.defaultdict(lambda: [0.0, 0])
for r in rows:
    if r is hdr or len(r) <= iv or r[iv] in ("Metric Value", ""):
        continue
    try:
        v = float(r[iv].replace(",", ""))
    except ValueError:
        continue
    unit = r[iu]
    us = v / 1000.0 if unit in ("ns", "nsecond") else (v * 1000.0 if unit in ("ms", "msecond") else v)
    name = re.sub(r"\(.*", "", r[ik]).replace("void ", "").replace("<unnamed>::", "")
    name = re.sub(r"<unnamed>::", "", name)[:110]
    agg[name][0] += us
    agg[name][1] += 1
tot = sum(v[0] for v in agg.values()) or 1.0
print(f"# {sum(v[1] for v in agg.values())} launches, {tot / 1000:.3f} ms serialised")
print("| ms | share | launches | kernel |\n|---|---|---|---|")
for name, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"| {us / 1000:.3f} | {100 * us / tot:.1f}% | {n} | `{name}` |")
