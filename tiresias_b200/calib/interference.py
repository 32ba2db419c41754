"""Co-location slowdown on one B200: the simulator packs up to three tasks on a device
(infra/device.py:53-55) and the reference stretches co-located tasks by a constant 1.2x
(infra/interference.py:1, infra/node.py:202 — which never fires). Here 1, 2 or 3 single-GPU
training processes share one GPU (time-sliced contexts, no MPS); each reports its device-timed iteration
time, and the table of ``shared / solo`` ratios goes into the calibration file.

    python -m tiresias_b200.calib.interference --models vgg16,transformer --out interference.json
"""
from __future__ import annotations

import argparse
import itertools
import json
import os
import subprocess
import sys
import tempfile
import time
from typing import Dict, List, Optional, Sequence


def run_group(models: Sequence[str], gpu: int, batch: int, iters: int, mode: str, timeout_s: float = 600.0) -> List[Dict]:
    """Run one process per entry of ``models`` on the same GPU, timed regions started together."""
    tmp = tempfile.mkdtemp(prefix="tb200_interf_")
    go = os.path.join(tmp, "go")
    root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    env = dict(os.environ, PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", ""))
    procs, outs = [], []
    for i, m in enumerate(models):
        out = os.path.join(tmp, f"r{i}.json")
        outs.append(out)
        procs.append(subprocess.Popen([sys.executable, "-m", "tiresias_b200.calib.solo", "--model", m, "--batch",
                                       str(batch), "--iters", str(iters), "--mode", mode, "--gpu", str(gpu),
                                       "--go-file", go, "--out", out], cwd=root, env=env))
    t0 = time.time()
    while not all(os.path.exists(o + ".ready") for o in outs):
        if time.time() - t0 > timeout_s or any(p.poll() not in (None, 0) for p in procs):
            for p in procs:
                p.kill()
            raise RuntimeError("co-location group failed to start")
        time.sleep(0.05)
    open(go, "w").write("1")
    for p in procs:
        p.wait(timeout=timeout_s)
    return [json.load(open(o)) for o in outs]


def measure(models: Sequence[str], gpu: int = 0, batch: int = 32, iters: int = 60, mode: str = "ps",
            max_colocated: int = 3) -> Dict:
    solo = {m: run_group([m], gpu, batch, iters, mode)[0] for m in models}
    table = {}
    factors = []
    combos = []
    for k in range(2, max_colocated + 1):
        combos += list(itertools.combinations_with_replacement(models, k))
    for combo in combos:
        res = run_group(combo, gpu, batch, iters, mode)
        entry = []
        for m, r in zip(combo, res):
            f = r["iter_ms"] / solo[m]["iter_ms"]
            entry.append({"model": m, "iter_ms": round(r["iter_ms"], 3), "slowdown": round(f, 3)})
            factors.append((k := len(combo), f))
        table["+".join(combo)] = entry
    by_k: Dict[int, List[float]] = {}
    for k, f in factors:
        by_k.setdefault(k, []).append(f)
    mean_by_k = {str(k): round(sum(v) / len(v), 3) for k, v in by_k.items()}
    return {"solo_iter_ms": {m: round(r["iter_ms"], 3) for m, r in solo.items()},
            "peak_mem_bytes": {m: r["peak_mem_bytes"] for m, r in solo.items()}, "colocated": table,
            "slowdown_by_tasks_per_gpu": mean_by_k,
            "interference_factor": round(mean_by_k.get("2", 1.0) - 1.0, 3),
            "reference_factor": 0.2, "batch": batch, "mode": mode}


def main(argv: Optional[Sequence[str]] = None) -> None:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--models", default="vgg16,transformer")
    ap.add_argument("--gpu", type=int, default=0)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--iters", type=int, default=60)
    ap.add_argument("--mode", default="ps")
    ap.add_argument("--max-colocated", type=int, default=3)
    ap.add_argument("--out", default="interference.json")
    a = ap.parse_args(list(argv) if argv is not None else None)
    res = measure(a.models.split(","), a.gpu, a.batch, a.iters, a.mode, a.max_colocated)
    with open(a.out, "w") as fh:
        json.dump(res, fh, indent=1)
    print(json.dumps({k: v for k, v in res.items() if k != "colocated"}))


if __name__ == "__main__":
    main()
