"""Regenerates tests/golden/*: runs the UNMODIFIED reference (/root/reference) under a wrapper that
seeds numpy/python RNGs, on synthetic traces, and keeps its cluster.csv. Only needed when the
fixtures change; the committed files let the parity tests run where the reference is absent.

    python tests/make_golden.py
"""
import glob
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = os.environ.get("TIRESIAS_REFERENCE", "/root/reference")
GOLD = os.path.join(ROOT, "tests", "golden")

WRAPPER = """
import sys, runpy, random
import numpy as np
np.random.seed({seed}); random.seed({seed})
sys.argv = ['run_sim.py'] + {argv!r}
sys.path.insert(0, '.')
try:
    runpy.run_path('run_sim.py', run_name='__main__')
except SystemExit:
    pass
"""

# name, trace kwargs, cluster flags, scheme, schedule, num_queue
CASES = [
    ("t12_yarn_fifo", dict(num_jobs=12, seed=3, oversized=64), (2, 4, 4), "yarn", "fifo", 1),
    ("t64_yarn_fifo", dict(num_jobs=64, seed=1), (4, 8, 8), "yarn", "fifo", 1),
    ("t64_yarn_horus", dict(num_jobs=64, seed=1), (4, 8, 8), "yarn", "horus", 1),
    ("t64_horus_horus", dict(num_jobs=64, seed=1), (2, 2, 4), "horus", "horus", 1),
    ("t64_horusp_q3", dict(num_jobs=64, seed=1), (2, 2, 4), "horus+", "horus+", 3),
    ("t64_horusp_q4", dict(num_jobs=64, seed=2, gap=8000.0), (2, 2, 4), "horus+", "horus+", 4),
    ("t64_gandiva", dict(num_jobs=64, seed=1, gap=8000.0), (2, 2, 4), "gandiva", "gandiva", 1),
    ("t64_yarn_gandiva", dict(num_jobs=64, seed=4, gap=8000.0), (2, 2, 4), "yarn", "gandiva", 1),
    ("t200_horusp_q5", dict(num_jobs=200, seed=5, gap=6000.0, interactive_every=7), (2, 4, 8), "horus+", "horus+", 5),
]


def main():
    from tiresias_b200.sim.synth import live_trace
    os.makedirs(GOLD, exist_ok=True)
    for name, tkw, (sw, npsw, gpn), scheme, schedule, nq in CASES:
        work = tempfile.mkdtemp(prefix="refrun_")
        ref = os.path.join(work, "ref")
        shutil.copytree(REF, ref, ignore=shutil.ignore_patterns(".git"))
        os.makedirs(os.path.join(ref, "data"), exist_ok=True)
        trace = live_trace(os.path.join(ref, "data", "trace.csv"), **tkw)
        argv = ["--num_switch", str(sw), "--num_node_p_switch", str(npsw), "--num_gpu_p_node", str(gpn),
                "--scheme", scheme, "--schedule", schedule, "--num_queue", str(nq),
                "--trace_file", "data/trace.csv", "--log_path", "gold"]
        with open(os.path.join(ref, "_wrap.py"), "w") as fh:
            fh.write(WRAPPER.format(seed=1, argv=argv))
        res = subprocess.run([sys.executable, "_wrap.py"], cwd=ref, capture_output=True, text=True)
        out = glob.glob(os.path.join(ref, "log", "gold", "*", "cluster.csv"))
        if not out:
            print(name, "FAILED", res.stderr[-2000:])
            continue
        shutil.copy(out[0], os.path.join(GOLD, name + ".cluster.csv"))
        shutil.copy(trace, os.path.join(GOLD, name + ".trace.csv"))
        with open(os.path.join(GOLD, name + ".args"), "w") as fh:
            fh.write(" ".join(argv[:-4]))
        print(name, "ok", sum(1 for _ in open(out[0])), "rows")
        shutil.rmtree(work, ignore_errors=True)


if __name__ == "__main__":
    main()
