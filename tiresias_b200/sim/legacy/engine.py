"""Runs one event-driven simulation: trace -> JobRegistry / TCluster / TLog -> scheduler loop ->
summary (reference wiring: upstream Tiresias ``main``; in the reference fork that wiring was removed,
run_sim.py:1709-1739, which is why its Tiresias schedulers are unreachable — SURVEY §0).
"""
from __future__ import annotations

import math
import os
import random
import time
from typing import Dict, List, Optional

from .. import util
from ..trace import JobTraceReader, TraceRow, detect_schema, live_rows_as_tiresias, read_tiresias_trace
from . import gittins as gi
from .cluster import TCluster
from .jobs import JobRegistry
from .log import TLog
from .schedulers import EVENT_SCHEDULES, Sim, dispatch


def parse_queue_limit(text: Optional[str], num_queue: int) -> Optional[List[int]]:
    """``--queue_limit 3600,7200,18000`` -> [3600, 7200, 18000] (README.md:55-62)."""
    if not text:
        return None
    limits = [int(float(x)) for x in str(text).replace(";", ",").split(",") if x.strip()]
    if limits != sorted(limits):
        raise ValueError(f"queue_limit must ascend: {limits}")
    return limits


def load_rows(flags) -> List[dict]:
    path = flags.trace_file
    if not path or not os.path.exists(path):
        raise ValueError(f"trace file {path!r} does not exist")
    if detect_schema(path) == "tiresias":
        return read_tiresias_trace(path)
    reader = JobTraceReader(path)
    reader.prepare_jobs()
    return live_rows_as_tiresias(reader)


def apply_calibration(rows: List[dict], calibration: Optional[Dict]) -> None:
    """Re-ground job durations in measured B200 iteration times: ``duration = iterations *
    iter_s[model][gpus]`` when the calibration file knows the model (tiresias_b200.calib)."""
    table = (calibration or {}).get("iter_s")
    if not table:
        return
    for r in rows:
        per = table.get(r.get("model_name") or "")
        if per and r.get("iterations"):
            keys = sorted(int(k) for k in per)
            g = max([k for k in keys if k <= r["num_gpu"]], default=keys[0])
            r["duration"] = max(1, int(math.ceil(r["iterations"] * float(per[str(g)]))))


def build(flags, output_dir: str, calibration: Optional[Dict] = None, rows: Optional[List[dict]] = None) -> Sim:
    rows = rows if rows is not None else load_rows(flags)
    apply_calibration(rows, calibration)
    from .. import models as _models
    _models.set_mem_table(str(getattr(flags, "pack_mem_util", "unit") or "unit") == "table")
    limits = parse_queue_limit(getattr(flags, "queue_limit", None), flags.num_queue)
    num_queue = flags.num_queue
    if limits is not None and num_queue < len(limits) + 1:
        num_queue = len(limits) + 1
    if flags.schedule.startswith(("dlas", "multi-dlas")) and num_queue < 2 and limits is None:
        num_queue = 2                                     # README default: two queues, [3600]
    jobs = JobRegistry(num_queue=num_queue, queue_limit=limits,
                       gittins_delta=int(getattr(flags, "gittins_delta", 3250) or 3250))
    for r in rows:
        jobs.add_job(r)
    jobs.sort_all_jobs()
    jobs.prepare_job_start_events()
    if flags.schedule in ("gittins", "dlas-gpu-gittins", "shortest-expected"):
        jobs.job_dist_data = gi.parse_job_dist(getattr(flags, "job_dist", None), jobs.gittins_delta,
                                               [j.duration * j.num_gpu for j in jobs.job_list])
    seed = getattr(flags, "seed", None)
    cluster = TCluster(flags, jobs, rng=random.Random(seed) if seed is not None else None)
    spec = getattr(flags, "cluster_spec", None)
    if spec:
        from ..infra import Infrastructure
        path = Infrastructure._resolve_spec_path(spec)
        if path:
            import csv
            with open(path) as fh:
                for row in csv.DictReader(fh):
                    cluster.set_spec(int(row["num_switch"]), int(row["num_node_p_switch"]),
                                     int(row["num_gpu_p_node"]), int(row["num_cpu_p_node"]),
                                     int(row["mem_p_node"]))
    cluster.init_infra()
    log = TLog(output_dir, flags, cluster)
    log.init_log()
    return Sim(flags, jobs, cluster, log, print_fn=util.print_fn)


def summarize(sim: Sim, wall: float) -> Dict:
    rows = sim.log.finished_rows
    n = len(rows)
    jct = [r[7] for r in rows]
    return {"makespan": max((r[5] for r in rows), default=0), "finished_jobs": n,
            "avg_jct": sum(jct) / n if n else float("nan"),
            "avg_pending": sum(r[9] for r in rows) / n if n else float("nan"),
            "queued_at_end": sum(1 for j in sim.jobs.job_list if j.status != "END"),
            "wall_seconds": wall, "ticks": sim.events}


def run_event_simulation(flags, output_dir: str, calibration: Optional[Dict] = None,
                         rows: Optional[List[dict]] = None) -> Dict:
    if flags.schedule not in EVENT_SCHEDULES:
        raise NotImplementedError(f"unknown schedule {flags.schedule!r}")
    sim = build(flags, output_dir, calibration, rows)
    util.print_fn(sim.cluster.print_cluster_spec())
    util.print_fn("---------------------------------- Get %d TF jobs in total ----------------------------------"
                  % sim.jobs.total_jobs())
    t0 = time.time()
    try:
        dispatch(sim)
    finally:
        sim.log.close()
    wall = time.time() - t0
    s = summarize(sim, wall)
    util.print_fn("Makespan: %d, finished %d, avg JCT %.2f, avg pending %.2f"
                  % (s["makespan"], s["finished_jobs"], s["avg_jct"], s["avg_pending"]))
    util.print_fn("Total Time Taken in seconds: %d" % wall)
    return s


class _RowsReader(JobTraceReader):
    """A :class:`JobTraceReader` fed from README-schema rows (so tick policies can replay them)."""

    def __init__(self, rows: List[TraceRow]):
        import numpy as np
        self._rows = rows
        self._times = np.asarray([r.normalized_time for r in rows], dtype=np.float64)
        self._cursor = 0
        self.trace_df = None
        self.delta_time_offset = 0

    def prepare_jobs(self) -> None:
        pass


def tiresias_trace_as_live(path: str, gpu_memory_gib: float = 32.0, scale_factor: float = 0.5) -> JobTraceReader:
    """View a README-schema trace through the live schema: one GPU per worker, memory from the
    model-size heuristic (sim.models.estimate_gpu_utilization), fixed 50/80 % utilisation."""
    from ..models import estimate_gpu_utilization, get_model_size, is_cnn
    rows = []
    for r in sorted(read_tiresias_trace(path), key=lambda x: x["submit_time"]):
        frac = estimate_gpu_utilization(get_model_size(r["model_name"]), is_cnn(r["model_name"]),
                                        gpu_memory_gib, add_noise=False)
        peak = min(0.95, frac) * gpu_memory_gib * (1 << 30)
        rows.append(TraceRow(r["job_id"], normalized_time=float(r["submit_time"]),
                             minutes=r["duration"] / scale_factor, gpu_per_container=1,
                             used_gpus=r["num_gpu"], gpu_utilization_avg=50.0, gpu_utilization_max=80.0,
                             memory_max=peak, memory_avg=peak / 2, model_name=r["model_name"],
                             iterations=r["iterations"]))
    return _RowsReader(rows)
