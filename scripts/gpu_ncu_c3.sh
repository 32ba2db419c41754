#!/bin/bash
mkdir -p gpurun_out
KERNELS="c3_3" bash scripts/gpu_ncu.sh > /dev/null 2>&1
mv gpurun_out/prof_c3_3.raw.csv gpurun_out/prof_c3_3_pair.raw.csv; mv gpurun_out/prof_c3_3.top_sass.txt gpurun_out/prof_c3_3_pair.top_sass.txt
TB200_CONV_PAIR=0 KERNELS="c3_3" bash scripts/gpu_ncu.sh > /dev/null 2>&1
python - <<'PY'
import csv
for tag in ("c3_3_pair", "c3_3"):
    f = open(f"gpurun_out/prof_{tag}.raw.csv"); lines = [l for l in f if not l.startswith("==")]
    r = csv.reader(lines); hdr = next(r); rows = [x for x in r if len(x) >= len(hdr)]
    idx = {h: i for i, h in enumerate(hdr)}
    row = rows[-1]
    print(tag, row[idx["Kernel Name"]][:40])
    for k in hdr:
        if any(w in k for w in ("gpu__time_duration.sum", "pipe_tensor_cycles_active_realtime.avg.pct", "xbar2l1tex_read_bytes.sum", "dram__bytes_read.sum", "lts__t_sectors_srcunit_tex_op_read.sum", "smsp__warp_issue_stalled", "l1tex__m_xbar2l1tex_read_sectors")) and "per_second" not in k and "pct_of_peak" not in k.replace("pipe_tensor_cycles_active_realtime.avg.pct_of_peak", ""):
            print("   ", k, row[idx[k]])
PY
head -30 gpurun_out/prof_c3_3_pair.top_sass.txt
