"""Side-by-side run of the UNMODIFIED reference simulator and this one on the same trace.

BASELINE.json's metric is two-part: "avg JCT & makespan per scheduler" (the reference produces it through
``run_sim.py:1709-1756`` / ``execute.py:5-53``) and PS push/pull samples/s (the reference has no execution path).
``bench.py`` carries the first part as a ``sim`` block: the reference from ``baseline/_ref`` (an untouched copy of
/root/reference — it has no setup.py / pyproject.toml, so ``pip install --target baseline/_ref`` refuses it, see
DESIGN.md section 2) is driven through its own ``run_sim.py`` under a wrapper that only seeds the two global RNGs
(its ``avg_gpu_utilization`` column is an unseeded ``np.random.normal`` draw), then this
simulator runs on the same trace and seed and the two ``cluster.csv`` files are compared byte for byte.
"""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
import sys
import tempfile
import time
from typing import Dict, List, Optional

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF_DIR = os.path.join(ROOT, "baseline", "_ref")

_WRAPPER = """
import sys, runpy, random, os
import numpy as np
np.random.seed({seed}); random.seed({seed})
os.chdir({ref!r})
sys.argv = ['run_sim.py'] + {argv!r}
sys.path.insert(0, {ref!r})
try:
    runpy.run_path(os.path.join({ref!r}, 'run_sim.py'), run_name='__main__')
except SystemExit:
    pass
"""

# BASELINE.json config 1: cluster_spec 4 switches x 32 nodes x 8 GPUs, yarn placement + fifo
DEFAULT_CASES = [
    {"name": "yarn_fifo_4x32x8", "scheme": "yarn", "schedule": "fifo", "num_queue": 1, "num_jobs": 64, "seed": 1,
     "cluster": (4, 32, 8)},
    {"name": "horus+_q3_4x32x8", "scheme": "horus+", "schedule": "horus+", "num_queue": 3, "num_jobs": 64, "seed": 1,
     "cluster": (4, 32, 8)},
    # long enough for the engines, not interpreter start-up, to dominate (reference ~20 s)
    {"name": "horus_300_4x32x8", "scheme": "horus", "schedule": "horus", "num_queue": 1, "num_jobs": 300, "seed": 2,
     "cluster": (4, 32, 8)},
]


def reference_available() -> Optional[str]:
    """None when ``baseline/_ref/run_sim.py`` can be run, else the reason."""
    if not os.path.isfile(os.path.join(REF_DIR, "run_sim.py")):
        return "baseline/_ref/run_sim.py not found (copy /root/reference to baseline/_ref)"
    try:
        import pandas  # noqa: F401  (the reference's trace reader)
    except Exception as e:  # noqa: BLE001
        return f"pandas missing: {e}"
    return None


def _argv(case: Dict, trace: str, log_path: str) -> List[str]:
    sw, npsw, gpn = case["cluster"]
    return ["--num_switch", str(sw), "--num_node_p_switch", str(npsw), "--num_gpu_p_node", str(gpn),
            "--scheme", case["scheme"], "--schedule", case["schedule"], "--num_queue", str(case["num_queue"]),
            "--trace_file", trace, "--log_path", log_path]


def _makespan(cluster_csv: bytes) -> Optional[int]:
    rows = [r for r in cluster_csv.split(b"\r\n") if r]
    if len(rows) < 2:
        return None
    try:
        return int(float(rows[-1].split(b",")[0]))
    except ValueError:
        return None


def run_reference(case: Dict, trace: str, seed: int = 1, timeout: float = 600.0) -> Dict:
    """The reference's own ``run_sim.py`` (cwd = baseline/_ref) on ``trace``; returns wall seconds, the bytes
    of its ``cluster.csv`` and the makespan (last ``delta``). Its log directory is removed afterwards."""
    log_path = f"bench_{os.getpid()}_{case['name'].replace('+', 'p')}"
    work = tempfile.mkdtemp(prefix="refsim_")
    wrapper = os.path.join(work, "_wrap.py")
    with open(wrapper, "w") as fh:
        fh.write(_WRAPPER.format(seed=seed, ref=REF_DIR, argv=_argv(case, trace, log_path)))
    t0 = time.perf_counter()
    res = subprocess.run([sys.executable, wrapper], capture_output=True, text=True, timeout=timeout)
    wall = time.perf_counter() - t0
    out = glob.glob(os.path.join(REF_DIR, "log", log_path, "*", "cluster.csv"))
    data = open(out[0], "rb").read() if out else b""
    shutil.rmtree(os.path.join(REF_DIR, "log", log_path), ignore_errors=True)
    try:
        os.rmdir(os.path.join(REF_DIR, "log"))           # leave the copy as it was (only if now empty)
    except OSError:
        pass
    shutil.rmtree(work, ignore_errors=True)
    if not data:
        return {"error": (res.stderr or "no cluster.csv written")[-300:], "wall_s": round(wall, 3)}
    return {"wall_s": round(wall, 3), "cluster_csv": data, "makespan": _makespan(data), "rows": data.count(b"\n") - 1}


def run_ours(case: Dict, trace: str, seed: int = 1) -> Dict:
    """This simulator in parity mode on the same trace / seed, in a subprocess (same cost model as the
    reference arm: interpreter + imports included in the wall time)."""
    work = tempfile.mkdtemp(prefix="oursim_")
    out_dir = os.path.join(work, "out")
    code = ("import sys; sys.path.insert(0, %r)\n"
            "from tiresias_b200.cli.run_sim import run\n"
            "from tiresias_b200.sim.flags import define_simulator_flags\n"
            "f = define_simulator_flags(); f.parse(%r); run(f, output_dir=%r)\n"
            % (ROOT, _argv(case, trace, "bench")[:-2] + ["--seed", str(seed), "--quiet"], out_dir))
    t0 = time.perf_counter()
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    wall = time.perf_counter() - t0
    path = os.path.join(out_dir, "cluster.csv")
    data = open(path, "rb").read() if os.path.exists(path) else b""
    summary = {}
    sp = os.path.join(out_dir, "summary.json")
    if os.path.exists(sp):
        import json
        summary = json.load(open(sp))
    shutil.rmtree(work, ignore_errors=True)
    if not data:
        return {"error": (res.stderr or "no cluster.csv written")[-300:], "wall_s": round(wall, 3)}
    return {"wall_s": round(wall, 3), "cluster_csv": data, "makespan": _makespan(data), "rows": data.count(b"\n") - 1,
            "avg_jct": summary.get("avg_jct"), "finished_jobs": summary.get("finished_jobs")}


def compare(cases: Optional[List[Dict]] = None, ours: bool = True, reference: bool = True) -> Dict:
    """The ``sim`` block of the bench line: per case both wall times, both makespans and whether the two
    ``cluster.csv`` files are identical."""
    from ..sim.synth import live_trace
    cases = cases or DEFAULT_CASES
    why = reference_available() if reference else "not requested"
    out: Dict = {"cases": [], "reference": "baseline/_ref (unmodified copy of /root/reference, own run_sim.py, seeded wrapper)"}
    if reference and why:
        out["reference_unavailable"] = why
    work = tempfile.mkdtemp(prefix="simtrace_")
    try:
        for case in cases:
            trace = live_trace(os.path.join(work, case["name"].replace("+", "p") + ".csv"), num_jobs=case["num_jobs"],
                               seed=case["seed"])
            row: Dict = {"trace": f"synthetic live-schema, {case['num_jobs']} jobs, seed {case['seed']}",
                         "config": f"{case['scheme']}/{case['schedule']} q{case['num_queue']} on "
                                   f"{'x'.join(map(str, case['cluster']))}"}
            ref = run_reference(case, trace) if (reference and not why) else None
            mine = run_ours(case, trace) if ours else None
            if ref is not None:
                row["reference_wall_s"] = ref.get("wall_s")
                row["reference_makespan"] = ref.get("makespan")
                if "error" in ref:
                    row["reference_error"] = ref["error"]
            if mine is not None:
                row["ours_wall_s"] = mine.get("wall_s")
                row["ours_makespan"] = mine.get("makespan")
                row["ours_avg_jct"] = mine.get("avg_jct")
                if "error" in mine:
                    row["ours_error"] = mine["error"]
            if ref is not None and mine is not None and "cluster_csv" in ref and "cluster_csv" in mine:
                row["makespan_equal"] = ref["makespan"] == mine["makespan"]
                row["cluster_csv_identical"] = ref["cluster_csv"] == mine["cluster_csv"]
                row["cluster_csv_rows"] = ref["rows"]
                row["speedup_wall"] = round(ref["wall_s"] / mine["wall_s"], 2) if mine["wall_s"] else None
            out["cases"].append(row)
    finally:
        shutil.rmtree(work, ignore_errors=True)
    return out
