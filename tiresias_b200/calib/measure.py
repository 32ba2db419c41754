"""Per-model iteration / start-up / communication times at 1, 2, 4, 8 GPUs -> ``calibration.json``.

    python -m tiresias_b200.calib.measure --gpus 8 --models resnet50,vgg16,transformer --comm ps --out calibration.json

Runs every (model, GPU count) once through the same worker pool the replay uses (so start-up cost is the
one a replayed job pays), all submitted at t = 0 under fifo. Output keys, all consumed by the simulator
(``--calibration``):

``iter_s[model][gpus]``            seconds per training iteration (device queue drained, max over ranks)
``startup_s[model][gpus]``         seconds from dispatch to the first iteration (model build, IPC exchange)
``comm_s_per_iter[model][gpus]``   ``iter_s[g] - iter_s[1]``: what synchronous data-parallel training adds per
                                   iteration on NVSwitch — the measured replacement of the reference's
                                   ``(model_size/bandwidth + cross*latency) * 2`` (core/network/network_service.py:34-37)
``interference_factor``            merged in from calib.interference when given (``--interference file``)
"""
from __future__ import annotations

import argparse
import json
import os
import tempfile
from typing import Dict, Optional, Sequence

from ..exec.replay import calibration_from_records, replay


def measure(gpus: int, models: Sequence[str], comm: str, iters: int, batch: int, out_dir: str,
            loopback: bool = False, step_sleep: float = 0.0, repeats: int = 2) -> Dict:
    os.makedirs(out_dir, exist_ok=True)
    trace = os.path.join(out_dir, "calibration_trace.csv")
    sizes = [g for g in (1, 2, 4, 8) if g <= gpus]
    with open(trace, "w") as fh:
        fh.write("job_id,num_gpu,submit_time,iterations,model_name,duration,interval\n")
        jid = 0
        for rep in range(repeats):                      # every configuration `repeats` times; the fastest counts
            for g in reversed(sizes):                   # big jobs first: the small ones then fill the box
                for m in models:
                    fh.write(f"{jid},{g},0,{iters},{m},1,0\n")
                    jid += 1
    res = replay(trace, gpus, "fifo", comm, os.path.join(out_dir, "calib_run"), None, loopback, batch,
                 step_sleep=step_sleep)
    cal = calibration_from_records(res["records"], reduce="min")
    comm_tab: Dict[str, Dict[str, float]] = {}
    for m, per in cal["iter_s"].items():
        base = per.get("1")
        if base is None:
            continue
        comm_tab[m] = {g: max(0.0, v - base) for g, v in per.items()}
    cal.update({"comm_s_per_iter": comm_tab, "source": f"calib.measure gpus={gpus} comm={comm} batch={batch} iters={iters}"})
    if any(r.get("peak_mem_bytes") for r in res["records"]):
        cal["peak_mem_bytes"] = {r["model"]: r["peak_mem_bytes"] for r in res["records"] if r["num_gpu"] == 1}
    return cal


def main(argv: Optional[Sequence[str]] = None) -> None:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--gpus", type=int, default=8)
    ap.add_argument("--models", default="resnet50,vgg16,transformer")
    ap.add_argument("--comm", default="ps", choices=["ps", "ring", "nccl"])
    ap.add_argument("--iters", type=int, default=60)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--interference", default=None, help="json written by calib.interference to merge in")
    ap.add_argument("--out", default="calibration.json")
    ap.add_argument("--work", default=None)
    ap.add_argument("--loopback", action="store_true")
    ap.add_argument("--step-sleep", type=float, default=0.0)
    ap.add_argument("--repeats", type=int, default=2)
    a = ap.parse_args(list(argv) if argv is not None else None)
    work = a.work or tempfile.mkdtemp(prefix="tb200_calib_")
    cal = measure(a.gpus, a.models.split(","), a.comm, a.iters, a.batch, work, a.loopback, a.step_sleep, a.repeats)
    if a.interference and os.path.exists(a.interference):
        inter = json.load(open(a.interference))
        cal["interference_factor"] = inter.get("interference_factor")
        cal["slowdown_by_tasks_per_gpu"] = inter.get("slowdown_by_tasks_per_gpu")
        cal["interference"] = inter
    with open(a.out, "w") as fh:
        json.dump(cal, fh, indent=1)
    print(json.dumps({k: cal[k] for k in ("iter_s", "startup_s", "comm_s_per_iter") if k in cal}))


if __name__ == "__main__":
    main()
