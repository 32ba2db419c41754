"""Replay a trace on the real box and compare with the simulator (BASELINE.json config 5).

    python -m tiresias_b200.exec.replay --trace trace.csv --gpus 8 --schedule fifo --comm ring \
        --calibration calibration.json --out replay_out

1. starts the worker pool (one process per GPU, :mod:`tiresias_b200.exec.worker`) and a TCPStore,
2. lets :class:`tiresias_b200.exec.controller.Controller` run the trace in real time,
3. simulates the same trace with the event engine on a 1 x 1 x G cluster (``--scheme count``), job
   durations = ``startup + iterations * iteration time`` taken from the calibration file (or, with
   ``--self-calibrate``, from this replay's own per-(model, GPUs) averages),
4. writes ``replay.json`` with per-job executed vs simulated JCT and the error summary.

``--loopback`` runs the whole thing on CPU tensors over gloo (tests / machines without GPUs).
"""
from __future__ import annotations

import argparse
import datetime
import json
import os
import socket
import subprocess
import sys
import tempfile
import time
from typing import Dict, List, Optional, Sequence

import torch.distributed as dist

from ..sim.trace import read_tiresias_trace
from .controller import Controller

MS = 1000.0
SIM_SCHEDULE = {"fifo": "fjf", "sjf": "sjf", "shortest": "shortest", "dlas-gpu": "dlas-gpu", "gittins": "gittins",
                "gandiva": "gandiva-slice"}


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def start_pool(gpus: int, ctl_port: int, loopback: bool, with_nccl: bool, log_path: str) -> subprocess.Popen:
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={gpus}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), "-m", "tiresias_b200.exec.worker",
           "--ctl-port", str(ctl_port)]
    if loopback:
        cmd.append("--loopback")
    if with_nccl:
        cmd.append("--with-nccl")
    root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    env = dict(os.environ, PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", ""))
    return subprocess.Popen(cmd, cwd=root, env=env, stdout=open(log_path, "w"), stderr=subprocess.STDOUT)


def calibration_from_records(records: Sequence[Dict], reduce: str = "mean") -> Dict:
    """``iter_s`` / ``startup_s`` per (model, GPUs) from finished jobs. ``reduce="min"`` (calibration runs,
    which repeat every configuration) discards first-use effects such as library autotuning."""
    acc: Dict[str, Dict[str, List[float]]] = {}
    start: Dict[str, Dict[str, List[float]]] = {}
    for r in records:
        if r["status"] != "DONE" or not r["iter_ms"]:
            continue
        acc.setdefault(r["model"], {}).setdefault(str(r["num_gpu"]), []).append(r["iter_ms"] / MS)
        start.setdefault(r["model"], {}).setdefault(str(r["num_gpu"]), []).append(r["startup_s"] or 0.0)
    fold = (lambda v: min(v)) if reduce == "min" else (lambda v: sum(v) / len(v))
    table = lambda d: {m: {g: fold(v) for g, v in per.items()} for m, per in d.items()}  # noqa: E731
    cal = {"iter_s": table(acc), "startup_s": table(start), "source": "self"}
    cal.update(measured_profiles(records))
    return cal


def measured_profiles(records: Sequence[Dict]) -> Dict:
    """K7: per-model GPU utilisation (NVML, sampled while the job ran) and peak device memory, the two
    per-task quantities the reference takes from the trace and perturbs with a random draw
    (infra/device.py:30-36, model/model_factory.py:68-106). ``{"gpu_util": {model: {"avg", "max"}},
    "peak_mem_bytes": {model: bytes}}`` — only models that were measured appear."""
    util: Dict[str, Dict[str, List[float]]] = {}
    mem: Dict[str, int] = {}
    for r in records:
        if r.get("status") != "DONE":
            continue
        if r.get("gpu_util_avg") is not None:
            u = util.setdefault(r["model"], {"avg": [], "max": []})
            u["avg"].append(float(r["gpu_util_avg"]))
            u["max"].append(float(r.get("gpu_util_max") or r["gpu_util_avg"]))
        if r.get("peak_mem_bytes"):
            mem[r["model"]] = max(mem.get(r["model"], 0), int(r["peak_mem_bytes"]))
    out: Dict = {}
    if util:
        out["gpu_util"] = {m: {"avg": sum(u["avg"]) / len(u["avg"]), "max": max(u["max"])} for m, u in util.items()}
    if mem:
        out["peak_mem_bytes"] = mem
    return out


def _lookup(table: Dict, model: str, gpus: int, default: float) -> float:
    per = (table or {}).get(model)
    if not per:
        return default
    keys = sorted(int(k) for k in per)
    pick = max([k for k in keys if k <= gpus], default=keys[0])
    return float(per[str(pick)])


def simulate(rows: Sequence[Dict], records: Sequence[Dict], calibration: Dict, gpus: int, schedule: str,
             queue_limit_s: Sequence[float], out_dir: str, scheme: str = "buddy",
             gittins_delta_s: float = 2.0, job_dist_ms: Optional[Sequence[int]] = None, pack: bool = False,
             max_share: int = 3, slowdown: Optional[Sequence[float]] = None, quantum_s: float = 10.0) -> Dict[str, Dict]:
    """Event simulation in milliseconds on a 1 x 1 x ``gpus`` cluster. ``scheme="buddy"`` places jobs on
    aligned power-of-two GPU blocks like the worker pool does (non-preemptive schedules); the preemptive
    ones hand out GPUs by counting, as their reference loops do."""
    from ..sim.flags import define_simulator_flags
    from ..sim.legacy.engine import run_event_simulation
    by_id = {r["job_id"]: r for r in records}
    sim_rows = []
    for row in rows:
        rec = by_id[str(row["job_id"])]
        g = rec["num_gpu"]
        it = _lookup(calibration.get("iter_s"), rec["model"], g, 0.05)
        st = _lookup(calibration.get("startup_s"), rec["model"], g, 0.0)
        sim_rows.append({"job_id": row["job_id"], "num_gpu": g, "submit_time": int(round(rec["submit"] * MS)),
                         "iterations": row["iterations"], "model_name": rec["model"],
                         "duration": max(1, int(round((st + it * row["iterations"]) * MS))), "interval": 0})
    flags = define_simulator_flags()
    sim_schedule = "fifo-pack" if (pack and schedule == "fifo") else SIM_SCHEDULE[schedule]
    argv = ["--schedule", sim_schedule, "--scheme", scheme, "--num_switch", "1", "--num_node_p_switch", "1",
            "--num_gpu_p_node", str(gpus), "--quiet"]
    if sim_schedule == "fifo-pack":
        # co-located single-GPU jobs advance at the measured rate for their tenancy
        argv += ["--pack_max_share", str(max_share), "--pack_slowdown",
                 ",".join(str(v) for v in (slowdown or calibration.get("interference_table") or (1.0, 1.2, 1.2)))]
    if sim_schedule == "gandiva-slice":
        argv += ["--slice_quantum", str(int(round(quantum_s * MS)))]
    if schedule == "dlas-gpu":
        argv += ["--queue_limit", ",".join(str(int(q * MS)) for q in queue_limit_s)]
    if schedule == "gittins":
        # same direction (highest index first), re-evaluation period, ranking rules and duration samples as the
        # controller (which knows only estimated iteration times when it builds its table)
        argv += ["--gittins_desc", "--gittins_clean", "--gittins_delta", str(int(round(gittins_delta_s * MS)))]
        if job_dist_ms:
            os.makedirs(out_dir, exist_ok=True)
            dist_path = os.path.join(out_dir, "job_dist.csv")
            with open(dist_path, "w") as fh:
                fh.write("duration\n" + "".join(f"{int(v)}\n" for v in job_dist_ms))
            argv += ["--job_dist", dist_path]
    if schedule not in ("fifo",):
        # restart cost of a resumed job, measured in this very replay: start-up + shard snapshot
        pre = [r for r in records if r.get("preempt")]
        if pre:
            over = sum((r.get("startup_s") or 0.0) + (r.get("snapshot_s") or 0.0) for r in pre) / len(pre)
            argv += ["--resume_overhead", str(int(round(over * MS)))]
    flags.parse(argv)
    os.makedirs(out_dir, exist_ok=True)
    from ..cli.run_sim import _attach_logging
    from ..sim import util
    handlers = _attach_logging(out_dir, True)
    try:
        run_event_simulation(flags, out_dir, None, rows=sim_rows)
    finally:
        for h in handlers:
            util.logger.removeHandler(h)
            h.close()
    import csv
    with open(os.path.join(out_dir, "jobs.csv")) as fh:
        return {r["job_id"]: r for r in csv.DictReader(fh)}


def compare(records: Sequence[Dict], sim: Dict[str, Dict]) -> Dict:
    rows, errs = [], []
    for r in records:
        s = sim.get(str(r["job_id"]))
        if r["status"] != "DONE" or s is None:
            continue
        sim_jct = float(s["JCT"]) / MS
        err = (sim_jct - r["jct"]) / r["jct"] * 100.0
        errs.append(err)
        rows.append({"job_id": r["job_id"], "model": r["model"], "num_gpu": r["num_gpu"], "exec_jct_s": round(r["jct"], 3),
                     "sim_jct_s": round(sim_jct, 3), "error_pct": round(err, 2)})
    n = len(rows)
    if not n:
        return {"jobs": [], "n": 0}
    avg_exec = sum(x["exec_jct_s"] for x in rows) / n
    avg_sim = sum(x["sim_jct_s"] for x in rows) / n
    mk_exec = max(r["end"] for r in records if r["end"] is not None)
    mk_sim = max(float(s["end_time"]) for s in sim.values()) / MS
    return {"n": n, "avg_jct_exec_s": round(avg_exec, 3), "avg_jct_sim_s": round(avg_sim, 3),
            "avg_jct_error_pct": round((avg_sim - avg_exec) / avg_exec * 100.0, 2),
            "mean_abs_job_error_pct": round(sum(abs(e) for e in errs) / n, 2),
            "median_abs_job_error_pct": round(sorted(abs(e) for e in errs)[n // 2], 2),
            "jct_weighted_abs_error_pct": round(sum(abs(x["sim_jct_s"] - x["exec_jct_s"]) for x in rows)
                                                / sum(x["exec_jct_s"] for x in rows) * 100.0, 2),
            "makespan_exec_s": round(mk_exec, 3), "makespan_sim_s": round(mk_sim, 3),
            "makespan_error_pct": round((mk_sim - mk_exec) / mk_exec * 100.0, 2), "jobs": rows}


def _estimated_iter_s(calibration_path: Optional[str], step_sleep: float) -> float:
    """Iteration time the gittins controller assumes before anything has run: the mean of a given
    calibration's 1-GPU entries, else the loopback's sleep per step, else 50 ms."""
    if calibration_path and os.path.exists(calibration_path):
        with open(calibration_path) as fh:
            per = [v.get("1") for v in (json.load(fh).get("iter_s") or {}).values() if v.get("1")]
        if per:
            return sum(per) / len(per)
    return step_sleep if step_sleep > 0 else 0.05


def _iter_table(calibration_path: Optional[str]) -> Optional[Dict]:
    if calibration_path and os.path.exists(calibration_path):
        with open(calibration_path) as fh:
            return json.load(fh).get("iter_s")
    return None


def replay(trace: str, gpus: int, schedule: str = "fifo", comm: str = "ps", out: str = "replay_out",
           calibration_path: Optional[str] = None, loopback: bool = False, batch: int = 32, time_scale: float = 1.0,
           queue_limit_s: Sequence[float] = (20.0,), check_every: int = 10, step_sleep: float = 0.0,
           iters_scale: float = 1.0, graph: bool = True, gittins_delta_s: float = 2.0, pack: bool = False,
           max_share: int = 3, quantum_s: float = 10.0, interference: Optional[Sequence[float]] = None) -> Dict:
    os.makedirs(out, exist_ok=True)
    rows = read_tiresias_trace(trace)
    for r in rows:
        r["iterations"] = max(1, int(round(r["iterations"] * iters_scale)))
    port = _free_port()
    store = dist.TCPStore("127.0.0.1", port, is_master=True, timeout=datetime.timedelta(days=1),
                          wait_for_workers=False)
    pool = start_pool(gpus, port, loopback, comm == "nccl", os.path.join(out, "pool.log"))
    ckpt_dir = tempfile.mkdtemp(prefix="tb200_ckpt_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    log_lines: List[str] = []

    def log(msg: str) -> None:
        log_lines.append(msg)
        print(msg, flush=True)
    ctl = Controller(store, gpus, rows, policy=schedule, comm=comm, batch=batch, time_scale=time_scale,
                     ckpt_dir=ckpt_dir, queue_limit=queue_limit_s, check_every=check_every, step_sleep=step_sleep,
                     graph=graph, log=log, gittins_delta_s=gittins_delta_s,
                     est_iter_s=_estimated_iter_s(calibration_path, step_sleep),
                     iter_table=_iter_table(calibration_path), pack=pack, max_share=max_share, quantum_s=quantum_s,
                     interference=interference if loopback else None)
    try:
        ctl.wait_ready()
        t_pool = time.time()
        ctl.run()
        wall = time.time() - t_pool
    finally:
        ctl.shutdown()
        try:
            pool.wait(timeout=120)
        except subprocess.TimeoutExpired:
            pool.kill()
    records = ctl.records()
    if calibration_path:
        with open(calibration_path) as fh:
            calibration = json.load(fh)
    else:
        calibration = calibration_from_records(records)
    if pack and not calibration_path:
        # self-calibration under packing: a job's own iteration time already contains the sharing it saw, so the
        # solo time is taken from the fastest attempt of its (model, GPUs) class
        calibration = calibration_from_records(records, reduce="min")
    sim = simulate(rows, records, calibration, gpus, schedule, queue_limit_s, os.path.join(out, "sim"),
                   "buddy" if schedule in ("fifo", "gandiva") else "count", gittins_delta_s, ctl.gittins_samples_ms,
                   pack=pack, max_share=max_share, slowdown=interference or calibration.get("interference_table"),
                   quantum_s=quantum_s)
    result = {"trace": trace, "gpus": gpus, "schedule": schedule, "comm": comm, "loopback": loopback,
              "pack": pack, "max_share": max_share, "quantum_s": quantum_s,
              "queue_limit_s": list(queue_limit_s),
              "wall_s": round(wall, 2), "calibration_source": calibration.get("source", calibration_path),
              "records": records, "comparison": compare(records, sim)}
    with open(os.path.join(out, "replay.json"), "w") as fh:
        json.dump(result, fh, indent=1)
    with open(os.path.join(out, "controller.log"), "w") as fh:
        fh.write("\n".join(log_lines) + "\n")
    return result


def resimulate(out: str, calibration_path: Optional[str] = None, scheme: str = "buddy") -> Dict:
    """Recompute the simulated side of an existing ``replay.json`` (no GPU needed)."""
    with open(os.path.join(out, "replay.json")) as fh:
        res = json.load(fh)
    rows = [{"job_id": r["job_id"], "iterations": r["iterations"]} for r in res["records"]]
    if calibration_path:
        with open(calibration_path) as fh:
            calibration = json.load(fh)
    else:
        calibration = calibration_from_records(res["records"])
    sim = simulate(rows, res["records"], calibration, res["gpus"], res["schedule"],
                   res.get("queue_limit_s") or (20.0,), os.path.join(out, "sim_" + scheme), scheme)
    res["comparison"] = compare(res["records"], sim)
    res["sim_scheme"] = scheme
    with open(os.path.join(out, "replay.json"), "w") as fh:
        json.dump(res, fh, indent=1)
    return res


def main(argv: Optional[Sequence[str]] = None) -> None:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--resimulate", default=None, metavar="DIR", help="recompute the comparison of an existing run")
    ap.add_argument("--sim-scheme", default="buddy", choices=["buddy", "count"])
    ap.add_argument("--trace", required=False)
    ap.add_argument("--gpus", type=int, default=8)
    ap.add_argument("--schedule", default="fifo", choices=sorted(SIM_SCHEDULE))
    ap.add_argument("--comm", default="ps", choices=["ps", "ring", "nccl"])
    ap.add_argument("--gittins-delta", type=float, default=2.0, help="gittins: service quantum in GPU-seconds")
    ap.add_argument("--out", default="replay_out")
    ap.add_argument("--calibration", default=None)
    ap.add_argument("--loopback", action="store_true")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--time-scale", type=float, default=1.0)
    ap.add_argument("--iters-scale", type=float, default=1.0)
    ap.add_argument("--queue-limit", default="20", help="dlas-gpu thresholds in GPU-seconds, comma separated")
    ap.add_argument("--check-every", type=int, default=10)
    ap.add_argument("--step-sleep", type=float, default=0.0, help="loopback only: seconds per iteration")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--pack", action="store_true", help="let single-GPU jobs share a GPU (up to --max-share)")
    ap.add_argument("--max-share", type=int, default=3)
    ap.add_argument("--quantum", type=float, default=10.0, help="gandiva: time slice in seconds")
    ap.add_argument("--interference", default=None, help="slowdown per tenancy, comma separated (simulation side; "
                                                         "on the loopback backend also emulated by the workers)")
    a = ap.parse_args(list(argv) if argv is not None else None)
    if a.resimulate:
        res = resimulate(a.resimulate, a.calibration, a.sim_scheme)
        print(json.dumps({k: v for k, v in res["comparison"].items() if k != "jobs"}))
        return
    res = replay(a.trace, a.gpus, a.schedule, a.comm, a.out, a.calibration, a.loopback, a.batch, a.time_scale,
                 [float(x) for x in a.queue_limit.split(",")], a.check_every, a.step_sleep, a.iters_scale,
                 not a.no_graph, a.gittins_delta, a.pack, a.max_share, a.quantum,
                 [float(x) for x in a.interference.split(",")] if a.interference else None)
    c = res["comparison"]
    print(json.dumps({k: v for k, v in c.items() if k != "jobs"}))


if __name__ == "__main__":
    main()
