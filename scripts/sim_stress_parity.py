#!/usr/bin/env python
"""One-off parity / speed check on the BASELINE cluster (4 switches x 32 nodes x 8 GPUs): runs the UNMODIFIED
reference (seeded wrapper of tests/make_golden.py) and this simulator on the same 300-400 job traces for
the four live policies, compares cluster.csv byte for byte and prints both wall times."""
import glob, os, shutil, subprocess, sys, tempfile, time
sys.path.insert(0, "/root/repo")
sys.path.insert(0, "/root/repo/tests")
from make_golden import WRAPPER, REF
from tiresias_b200.sim.synth import live_trace
from tiresias_b200.cli.run_sim import run
from tiresias_b200.sim.flags import define_simulator_flags
CASES = [("yarn","fifo",1,dict(num_jobs=400, seed=11)), ("horus","horus",1,dict(num_jobs=300, seed=12)),
         ("horus+","horus+",3,dict(num_jobs=300, seed=13)), ("gandiva","gandiva",1,dict(num_jobs=300, seed=14, gap=3000.0))]
for scheme, schedule, nq, tkw in CASES:
    work = tempfile.mkdtemp(prefix="stress_")
    ref = os.path.join(work, "ref")
    shutil.copytree(REF, ref, ignore=shutil.ignore_patterns(".git"))
    os.makedirs(os.path.join(ref, "data"), exist_ok=True)
    trace = live_trace(os.path.join(ref, "data", "trace.csv"), **tkw)
    argv = ["--num_switch", "4", "--num_node_p_switch", "32", "--num_gpu_p_node", "8", "--scheme", scheme,
            "--schedule", schedule, "--num_queue", str(nq), "--trace_file", "data/trace.csv", "--log_path", "gold"]
    open(os.path.join(ref, "_wrap.py"), "w").write(WRAPPER.format(seed=1, argv=argv))
    t = time.time()
    res = subprocess.run([sys.executable, "_wrap.py"], cwd=ref, capture_output=True, text=True)
    t_ref = time.time() - t
    out = glob.glob(os.path.join(ref, "log", "gold", "*", "cluster.csv"))
    if not out:
        print(scheme, schedule, "REFERENCE FAILED", res.stderr[-500:]); continue
    want = open(out[0], "rb").read()
    f = define_simulator_flags()
    f.parse(argv[:-4] + ["--trace_file", trace, "--seed", "1", "--quiet"])
    t = time.time()
    run(f, output_dir=os.path.join(work, "ours"))
    t_ours = time.time() - t
    got = open(os.path.join(work, "ours", "cluster.csv"), "rb").read()
    same = got == want
    first = None
    if not same:
        g, w = got.split(b"\r\n"), want.split(b"\r\n")
        first = next((i for i, (a, b) in enumerate(zip(g, w)) if a != b), min(len(g), len(w)))
    print(scheme, schedule, nq, "rows", want.count(b"\n"), "identical", same, "first diff row", first,
          "ref %.1fs ours %.1fs" % (t_ref, t_ours), flush=True)
    shutil.rmtree(work, ignore_errors=True)
