"""CSV logs (reference: log_manager.py:4-130, C13; per-job row layout from the legacy log.py:316-330,
D13).

Files under the run directory: ``cluster.csv`` (one row per tick), ``job.csv`` (the reference's file
name; on the live path it only ever holds the header — SURVEY §0), ``jobs.csv`` (the README's name:
same header plus one row per finished job, written by this framework for every scheduler),
``cpu.csv``, ``gpu.csv``, ``memory.csv``, ``network.csv`` (headers; filled by the legacy/event
schedulers). Rows use the ``csv`` module's ``\\r\\n`` terminator and ``str()`` of every value, which
is what makes ``avg_gpu_utilization`` print as ``[4.51252041]`` when it came from a NumPy draw.
"""
from __future__ import annotations

import csv
import os
from typing import List

import numpy as np

CLUSTER_HEADER = ["delta", "num_idle_nodes", "num_busy_nodes", "num_busy_gpus", "num_idle_gpus",
                  "avg_gpu_utilization", "avg_gpu_memory_allocated", "avg_pending_time",
                  "median_pending_time", "max_pending_time", "num_running_jobs", "num_queuing_jobs",
                  "num_finish_jobs"]
JOB_HEADER = ["time", "job_id", "num_gpu", "submit_time", "start_time", "end_time", "executed_time",
              "JCT", "duration", "pending_time", "preempt", "promote"]
JOB_HEADER_COUNT = JOB_HEADER[:-1] + ["resume", "promote"]
GPU_DEMANDS_HEADER = ["time", "1-GPU", "2-GPU", "4-GPU", "8-GPU", "12-GPU", "16-GPU", "24-GPU", "32-GPU"]


class ArrayLike(float):
    """A float that prints the way a 1-element float64 ndarray does (``[4.51252041]``)."""

    def __str__(self) -> str:
        return str(np.array([float(self)]))

    __repr__ = __str__


class LogInfo:
    """One ``cluster.csv`` row (12 fields + the tick)."""

    __slots__ = ("idle_ns", "busy_ns", "busy_gs", "idle_gs", "avg_g_utils", "avg_g_mem", "avg_pending",
                 "median_pending", "max_pending", "num_running_jobs", "num_queuing_jobs",
                 "num_finish_jobs")

    def __init__(self, num_idle_nodes, num_busy_nodes, num_busy_gpus, num_idle_gpus,
                 avg_gpu_utilization, avg_gpu_memory_allocated, avg_pending_time, median_pending_time,
                 max_pending_time, num_running_jobs, num_queuing_jobs, num_finish_jobs) -> None:
        self.idle_ns, self.busy_ns = num_idle_nodes, num_busy_nodes
        self.busy_gs, self.idle_gs = num_busy_gpus, num_idle_gpus
        self.avg_g_utils, self.avg_g_mem = avg_gpu_utilization, avg_gpu_memory_allocated
        self.avg_pending, self.median_pending = avg_pending_time, median_pending_time
        self.max_pending = max_pending_time
        self.num_running_jobs, self.num_queuing_jobs = num_running_jobs, num_queuing_jobs
        self.num_finish_jobs = num_finish_jobs

    def as_row(self, delta) -> list:
        return [delta, self.idle_ns, self.busy_ns, self.busy_gs, self.idle_gs, self.avg_g_utils,
                self.avg_g_mem, self.avg_pending, self.median_pending, self.max_pending,
                self.num_running_jobs, self.num_queuing_jobs, self.num_finish_jobs]


class LogManager:
    def __init__(self, log_path: str, flags):
        self.log_path = log_path
        self.flags = flags
        self.is_count = flags.scheme == "count"
        self.cluster_stats_header = list(CLUSTER_HEADER)
        self._cluster_fh = None
        self._cluster_writer = None
        self._jobs_fh = None
        self._jobs_writer = None
        self.flush_every = 1 if getattr(flags, "flush_stdout", True) else 256
        self._since_flush = 0

    # -- set-up ------------------------------------------------------------------------------------
    def init(self, infrastructure) -> None:
        p = self.log_path
        self.log_cluster = os.path.join(p, "cluster.csv")
        self.log_job = os.path.join(p, "job.csv")
        self.log_jobs = os.path.join(p, "jobs.csv")
        if not self.is_count:
            self.log_cpu = os.path.join(p, "cpu.csv")
            self.log_gpu = os.path.join(p, "gpu.csv")
            self.log_network = os.path.join(p, "network.csv")
            self.log_mem = os.path.join(p, "memory.csv")
        self._init_all_csv(infrastructure)

    @staticmethod
    def _write_header(path: str, header: List[str]) -> None:
        with open(path, "w+", newline="") as fh:
            csv.writer(fh).writerow(header)

    def job_header(self) -> List[str]:
        if self.flags.schedule == "gpu-demands":
            return list(GPU_DEMANDS_HEADER)
        return list(JOB_HEADER_COUNT if self.is_count else JOB_HEADER)

    def _init_all_csv(self, infrastructure) -> None:
        self._write_header(self.log_cluster, self.cluster_stats_header)
        if not self.is_count:
            n_nodes = len(infrastructure.nodes)
            self._write_header(self.log_cpu, ["time"] + ["cpu" + str(i) for i in range(n_nodes)])
            self._write_header(self.log_gpu,
                               ["time"] + ["gpu" + str(i) for i in range(infrastructure.get_total_gpus())])
            self._write_header(self.log_mem, ["time", "max", "99th", "95th", "med"])
            titles = ["time"]
            for i in range(n_nodes):
                titles += ["in" + str(i), "out" + str(i)]
            self._write_header(self.log_network, titles)
        self._write_header(self.log_job, self.job_header())
        self._write_header(self.log_jobs, self.job_header())
        assert os.path.exists(self.log_cluster)
        self._cluster_fh = open(self.log_cluster, "a+", newline="")
        self._cluster_writer = csv.writer(self._cluster_fh)
        self._jobs_fh = open(self.log_jobs, "a+", newline="")
        self._jobs_writer = csv.writer(self._jobs_fh)

    # -- rows ----------------------------------------------------------------------------------------
    def step_cluster(self, loginfo: LogInfo, delta) -> None:
        self._cluster_writer.writerow(loginfo.as_row(delta))
        self._since_flush += 1
        if self._since_flush >= self.flush_every:
            self._cluster_fh.flush()
            self._since_flush = 0

    def job_complete(self, row: list, also_job_csv: bool = False) -> None:
        """Append one per-job row to ``jobs.csv`` (and to ``job.csv`` for the event-driven
        schedulers, whose reference counterpart writes rows there)."""
        self._jobs_writer.writerow(row)
        self._jobs_fh.flush()
        if also_job_csv:
            with open(self.log_job, "a+", newline="") as fh:
                csv.writer(fh).writerow(row)

    def append(self, which: str, row: list) -> None:
        """Append a row to one of cpu / gpu / memory / network .csv."""
        path = {"cpu": "log_cpu", "gpu": "log_gpu", "memory": "log_mem", "network": "log_network"}[which]
        with open(getattr(self, path), "a+", newline="") as fh:
            csv.writer(fh).writerow(row)

    def close(self) -> None:
        for fh in (self._cluster_fh, self._jobs_fh):
            if fh is not None:
                fh.close()
        self._cluster_fh = self._jobs_fh = None
