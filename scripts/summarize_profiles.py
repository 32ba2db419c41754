#!/usr/bin/env python
"""Turns gpurun_out/prof_*.ncu-rep (+ kernels.log, launches_step.csv) into the tracked summaries under
profiles/: ncu_summary_<tag>.md (one row per captured launch), sass_evidence_<tag>.txt (tcgen05 / TMA /
peer-memory mnemonics per kernel from cuobjdump), kernels_<tag>.json, step_launches_<tag>.md.

    python scripts/summarize_profiles.py r1a
"""
import collections
import csv
import glob
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
GP = os.path.join(ROOT, "gpurun_out")
METRICS = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
           ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
           ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %"),
           ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps %"),
           ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block")]
SASS = re.compile(r"\b(UTC[A-Z]*MMA|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|UTCBAR|SYNCS|HMMA|LDG|STG|LD|ST|RED|ATOM|ATOMG|MEMBAR|"
                  r"ERRBAR|FENCE|UTCCP)\b((?:\.\w+)*)")


def ncu_rows(rep):
    if rep.endswith(".raw.csv"):
        text = open(rep).read()
    else:
        text = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(text)))
    if len(rows) < 3:
        return []
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        d = {"kernel": re.sub(r"\(.*", "", name).replace("void ", "").replace("<unnamed>::", "")}
        for m, short in METRICS:
            if m in hdr:
                v = r[hdr.index(m)]
                try:
                    v = f"{float(v.replace(',', '')):.4g}"
                except ValueError:
                    pass
                d[short] = (v + " " + units[hdr.index(m)]).strip()
        out.append(d)
    return out


def write_ncu(tag):
    lines = [f"# ncu --set full captures ({tag}; --clock-control none; 3 launches per kernel after warm-up)", "",
             "| capture | kernel | " + " | ".join(s for _, s in METRICS) + " |", "|---|---|" + "---|" * len(METRICS)]
    for rep in sorted(glob.glob(os.path.join(GP, "prof_*.raw.csv")) or glob.glob(os.path.join(GP, "prof_*.ncu-rep"))):
        name = os.path.basename(rep)[5:].replace(".raw.csv", "").replace(".ncu-rep", "")
        for d in ncu_rows(rep):
            lines.append(f"| {name} | `{d['kernel'][:60]}` | " + " | ".join(d.get(s, "") for _, s in METRICS) + " |")
    open(os.path.join(OUT, f"ncu_summary_{tag}.md"), "w").write("\n".join(lines) + "\n")
    tops = sorted(glob.glob(os.path.join(GP, "prof_*.top_sass.txt")))
    if tops:
        with open(os.path.join(OUT, f"ncu_hot_sass_{tag}.txt"), "w") as fh:
            fh.write("Most-sampled SASS instructions per captured kernel (ncu --page source, warp stall sampling)\n\n")
            for t in tops:
                fh.write(f"#### {os.path.basename(t)[5:-13]}\n" + open(t).read() + "\n")


def write_sass(tag):
    so = os.path.join(ROOT, "tiresias_b200", "lib", "libtiresias_b200.so")
    sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    per = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            per[cur] = collections.Counter()
            continue
        if cur and "/*" in line:
            body = line.split("*/", 1)[-1]
            for mm in SASS.finditer(body):
                mods = [x for x in mm.group(2).split(".") if x in ("SYS", "STRONG", "GPU", "E", "2D", "TILED", "128x256b", "32x32b", "ARRIVE", "TRANS64")]
                key = mm.group(1) + ("." + ".".join(mods) if mods else "")
                if mm.group(1) in ("LD", "ST", "LDG", "STG") and "SYS" not in mods:
                    key = mm.group(1)
                per[cur][key] += 1
    with open(os.path.join(OUT, f"sass_evidence_{tag}.txt"), "w") as fh:
        fh.write("cuobjdump -sass tiresias_b200/lib/libtiresias_b200.so : instruction-class counts per kernel\n"
                 "UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG / UTMASTG = TMA load / store, SYNCS = mbarrier,\n"
                 "*.SYS = system-scope accesses (flags and data on peer-mapped NVLink pointers in the push / ring kernels)\n\n")
        for fn, cnt in per.items():
            dem = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip()
            dem = re.sub(r"\(.*", "", dem.replace("(anonymous namespace)::", "").replace("void ", ""))[:100]
            fh.write(f"{dem}\n    " + ", ".join(f"{k} x{v}" for k, v in sorted(cnt.items())) + "\n")


def write_kernels(tag):
    klog = os.path.join(GP, "kernels.log")
    if os.path.exists(klog):
        rows = [json.loads(l) for l in open(klog) if l.startswith("{")]
        json.dump(rows, open(os.path.join(OUT, f"kernels_{tag}.json"), "w"), indent=1)


def write_step(tag):
    ls = os.path.join(GP, "launches_step.csv")
    if not os.path.exists(ls):
        return
    txt = "\n".join(l for l in open(ls).read().splitlines() if not l.startswith("=="))
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot = 0.0
    for row in csv.DictReader(io.StringIO(txt)):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", "")) * {"ns": 1, "us": 1e3, "ms": 1e6}.get(row["Metric Unit"], 1)
        k = re.sub(r"\(.*", "", row["Kernel Name"])[:100]
        agg[k][0] += 1
        agg[k][1] += v
        tot += v
    with open(os.path.join(OUT, f"step_launches_{tag}.md"), "w") as fh:
        fh.write(f"# One eager ResNet-50 PS step (batch 64, 1 GPU) under ncu: {sum(a[0] for a in agg.values())} "
                 f"launches, {tot / 1e6:.2f} ms serialised (cold cache; compare shares)\n\n"
                 "| ms | share | launches | kernel |\n|---|---|---|---|\n")
        for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:45]:
            fh.write(f"| {t / 1e6:.3f} | {100 * t / tot:.1f}% | {n} | `{k}` |\n")


def main():
    os.makedirs(OUT, exist_ok=True)
    tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
    write_ncu(tag)
    write_sass(tag)
    write_kernels(tag)
    write_step(tag)


if __name__ == "__main__":
    main()
