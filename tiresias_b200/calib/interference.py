"""Co-location slowdown on one B200: the simulator packs up to three tasks on a device
(infra/device.py:53-55) and the reference stretches co-located tasks by a constant 1.2x
(infra/interference.py:1, infra/node.py:202 — which never fires). Here 1, 2 or 3 single-GPU
training jobs share one GPU and each reports its device-timed iteration time; the table of ``shared / solo``
ratios goes into the calibration file (``interference_table`` = mean slowdown at 1, 2, 3 tenants, what the
``fifo-pack`` simulation and the executor's packing mode use).

Two sharing mechanisms:

``--share streams`` (default)  one process, one CUDA context, every job a thread with its own streams and its own
                               captured CUDA graph — the mechanism the worker pool's packing mode uses
                               (exec/worker.py). Kernels of different jobs really overlap on the SMs, so the
                               slowdown is what contention for SMs, L2 and HBM costs: typically well under k
                               for k tenants when the jobs alone do not fill the GPU.
``--share processes``          separate processes without MPS: contexts are time-sliced, the slowdown is ~k by
                               construction (round 1's measurement: 2.19x / 3.25x).

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


def run_group_streams(models: Sequence[str], gpu: int, batch: int, iters: int, mode: str) -> List[Dict]:
    """All jobs inside THIS process: one thread + one Trainer (own streams, own CUDA graph) per job, timed
    regions released together by a barrier. Same context => their kernels run concurrently."""
    import threading

    import torch

    from ..models import zoo
    from ..parallel.trainer import Trainer
    torch.cuda.set_device(gpu)
    dev = torch.device("cuda", gpu)
    n = len(models)
    ready, go = threading.Barrier(n), threading.Barrier(n)
    setup = threading.Lock()                    # model build, allocator growth and graph capture one tenant at a time
    out: List[Optional[Dict]] = [None] * n
    errs: List[str] = []

    def body(i: int, name: str) -> None:
        try:
            torch.cuda.set_device(gpu)
            with setup:
                spec = zoo.resolve(name)
                tr = Trainer(spec, batch, mode=mode, use_graph=True, device=dev)
                x, y = zoo.synthetic_batch(spec, batch, seed=i)
                for _ in range(6):              # allocator growth, graph capture (eager steps serialise on the
                    tr.step(x, y)               # autograd thread; the timed replays below do not)
                tr.loss()
            ready.wait()
            start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            go.wait()
            with torch.cuda.stream(tr.stream):
                start.record()
            for _ in range(iters):
                tr.step_resident()
            with torch.cuda.stream(tr.stream):
                end.record()
            tr.synchronize()
            out[i] = {"model": name, "iters": iters, "iter_ms": start.elapsed_time(end) / iters,
                      "peak_mem_bytes": int(torch.cuda.max_memory_allocated(dev))}
            tr.close()
        except Exception as e:  # noqa: BLE001
            errs.append(f"{name}: {type(e).__name__}: {e}")
            for b in (ready, go):
                b.abort()

    threads = [threading.Thread(target=body, args=(i, m)) for i, m in enumerate(models)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if errs or any(o is None for o in out):
        raise RuntimeError("co-location group failed: " + "; ".join(errs))
    return out  # type: ignore[return-value]


def measure(models: Sequence[str], gpu: int = 0, batch: int = 32, iters: int = 60, mode: str = "ps",
            max_colocated: int = 3, share: str = "streams") -> Dict:
    if share == "streams":
        group = lambda ms: run_group_streams(ms, gpu, batch, iters, mode)          # noqa: E731
    else:
        group = lambda ms: run_group(ms, gpu, batch, iters, mode)                  # noqa: E731
    run_group_ = group
    solo = {m: run_group_([m])[0] for m in models}
    table = {}
    factors = []
    combos = []
    for k in range(2, max_colocated + 1):
        combos += list(itertools.combinations_with_replacement(models, k))
    for combo in combos:
        res = run_group_(combo)
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
            "interference_table": [1.0] + [mean_by_k.get(str(k), float(k)) for k in range(2, max_colocated + 1)],
            "interference_factor": round(mean_by_k.get("2", 1.0) - 1.0, 3),
            "reference_factor": 0.2, "batch": batch, "mode": mode, "share": share}


def main(argv: Optional[Sequence[str]] = None) -> None:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--models", default="vgg16,transformer")
    ap.add_argument("--gpu", type=int, default=0)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--iters", type=int, default=60)
    ap.add_argument("--mode", default="ps")
    ap.add_argument("--max-colocated", type=int, default=3)
    ap.add_argument("--share", default="streams", choices=["streams", "processes"])
    ap.add_argument("--out", default="interference.json")
    a = ap.parse_args(list(argv) if argv is not None else None)
    res = measure(a.models.split(","), a.gpu, a.batch, a.iters, a.mode, a.max_colocated, a.share)
    with open(a.out, "w") as fh:
        json.dump(res, fh, indent=1)
    print(json.dumps({k: v for k, v in res.items() if k != "colocated"}))


if __name__ == "__main__":
    main()
