"""Job registry of the event-driven (Tiresias) schedulers.

The reference's schedulers (run_sim.py:98-1707) talk to a global ``JOBS`` object whose module
(upstream ``jobs.py``) is absent from the reference; its contract is reconstructed from usage
(SURVEY §2.3 "implied legacy API"). :class:`TJob` carries the per-job keys the schedulers read and
write (``job['num_gpu']``-style access works too), :class:`JobRegistry` the lists, queues, MLFQ
thresholds, Gittins table and per-GPU-count groups.

Network bookkeeping follows the one-worker-one-PS-shard-per-GPU layout of infra/switch.py:97-107:
``w_network[i]`` = whole model (MB) per worker, ``ps_network[s]`` = sum of the tensors sharded
round-robin onto PS ``s`` (core/models.py tensor tables).
"""
from __future__ import annotations

import math
import sys
from typing import Dict, List, Optional

from ..models import get_model_with_scale, per_worker_mem, ps_mem, worker_mem

MAXSIZE = sys.maxsize
README_QUEUE_LIMITS = {2: [3600], 4: [3600, 7200, 18000]}      # README.md:55-62


class TJob:
    """One trace job. Times are integer seconds."""

    __slots__ = ("job_idx", "job_id", "num_gpu", "submit_time", "r_submit_time", "duration", "iterations",
                 "model_name", "interval", "status", "start_time", "end_time", "last_check_time",
                 "total_executed_time", "executed_time", "pending_time", "last_pending_time",
                 "remaining_time", "remaining_gputime", "remaining_expected", "q_id", "rank", "preempt",
                 "resume", "promote", "placements", "gpus", "model", "ps_network", "w_network",
                 "model_scale", "last_start_time", "progress", "slice_start")

    def __init__(self, row: dict, idx: int):
        self.job_idx = idx
        self.job_id = row["job_id"]
        self.num_gpu = int(row["num_gpu"])
        self.submit_time = int(row["submit_time"])
        self.r_submit_time = -self.submit_time
        self.duration = int(float(row["duration"]))
        self.iterations = int(row.get("iterations") or 0)
        self.model_name = row.get("model_name") or ""
        self.interval = int(row.get("interval") or 0)
        self.model_scale = int(row.get("model_scale") or 1)
        self.status = "ADDED"
        self.start_time = MAXSIZE
        self.end_time = 0
        self.last_start_time = 0
        self.last_check_time = 0
        self.total_executed_time = 0
        self.executed_time = 0
        self.pending_time = 0
        self.last_pending_time = 0
        self.remaining_time = self.duration
        self.remaining_gputime = self.duration * self.num_gpu
        self.remaining_expected = 0
        self.q_id = 0
        self.rank = MAXSIZE
        self.preempt = self.resume = self.promote = 0
        self.placements: List[dict] = []
        self.gpus: List[int] = []
        self.model = get_model_with_scale(self.model_name, self.model_scale)
        self._network_load()

    def _network_load(self) -> None:
        n = self.num_gpu
        if n == 1:
            self.ps_network, self.w_network = [], [0]
            return
        self.w_network = [self.model["total_size"]] * n
        shards = [0.0] * n
        for i, t in enumerate(self.model["tensors"]):
            shards[i % n] += t
        self.ps_network = [round(s, 1) for s in shards]

    # dict-style access, as the reference's job dicts
    def __getitem__(self, key):
        return getattr(self, key)

    def __setitem__(self, key, value) -> None:
        setattr(self, key, value)

    def __contains__(self, key) -> bool:
        return key in self.__slots__

    def __repr__(self) -> str:
        return f"TJob({self.job_id}, {self.num_gpu}G, submit={self.submit_time}, dur={self.duration}, {self.status})"


class GpuJobGroup:
    """Jobs of one GPU-count class with their own MLFQ and GPU reservation (multi-dlas-gpu)."""

    def __init__(self, num_gpu: int, num_queue: int, queue_limit: List[int]):
        self.num_gpu = num_gpu
        self.name = f"{num_gpu}-GPU"
        self.num_queue = num_queue
        self.queue_limit = list(queue_limit)
        self.queues: List[List[TJob]] = [[] for _ in range(num_queue)]
        self.runnable_jobs: List[TJob] = []
        self.total_gpu = 0
        self.free_gpu = 0
        self.total_job = 0
        self.end_job = 0

    def empty_gpu_alloc(self) -> None:
        self.free_gpu = self.total_gpu

    def alloc_free_gpus(self, need: int) -> bool:
        if self.free_gpu >= need:
            self.free_gpu -= need
            return True
        return False

    def get_gpu_demands(self) -> int:
        return self.num_gpu * len(self.runnable_jobs)

    def set_reservation(self, reserved: int) -> None:
        used = self.total_gpu - self.free_gpu
        self.total_gpu = reserved
        self.free_gpu = max(0, reserved - used)


class JobRegistry:
    """The ``JOBS`` object."""

    def __init__(self, num_queue: int = 1, queue_limit: Optional[List[int]] = None,
                 gittins_delta: int = 3250):
        self.job_list: List[TJob] = []
        self.job_events: List[dict] = []
        self.pending_jobs: List[TJob] = []
        self.running_jobs: List[TJob] = []
        self.runnable_jobs: List[TJob] = []
        self.completed_jobs: List[TJob] = []
        self.num_queue = max(1, int(num_queue))
        if queue_limit is None:
            queue_limit = README_QUEUE_LIMITS.get(self.num_queue, [3600 * (2 ** i) for i in range(self.num_queue - 1)])
        self.queue_limit = list(queue_limit)
        if len(self.queue_limit) < self.num_queue - 1:
            raise ValueError(f"{self.num_queue} queues need {self.num_queue - 1} thresholds, got {self.queue_limit}")
        self.queues: List[List[TJob]] = [[] for _ in range(self.num_queue)]
        self.gittins_delta = gittins_delta
        self.job_dist_data: Optional[dict] = None
        self.gpu_job: Dict[int, object] = {}
        self.worker_mem, self.ps_mem, self.p_w_mem = worker_mem, ps_mem, per_worker_mem
        self.multi_dlas = False

    # -- loading --------------------------------------------------------------------------------------
    def add_job(self, row: dict) -> TJob:
        job = TJob(row, len(self.job_list))
        self.job_list.append(job)
        return job

    def sort_all_jobs(self) -> None:
        self.job_list.sort(key=lambda j: j.submit_time)
        for i, j in enumerate(self.job_list):
            j.job_idx = i

    def prepare_job_start_events(self) -> None:
        by_time: Dict[int, dict] = {}
        for job in self.job_list:
            ev = by_time.get(job.submit_time)
            if ev is None:
                ev = by_time[job.submit_time] = {"time": job.submit_time, "start_jobs": [], "end_jobs": []}
                self.job_events.append(ev)
            ev["start_jobs"].append(job)
            job.status = "EVENT"
        self.job_events.sort(key=lambda e: e["time"])

    def add_job_end_event(self, job: TJob) -> None:
        for ev in self.job_events:
            if ev["time"] == job.end_time:
                ev["end_jobs"].append(job)
                break
        else:
            self.job_events.append({"time": job.end_time, "start_jobs": [], "end_jobs": [job]})
        self.job_events.sort(key=lambda e: e["time"])

    # -- state moves ---------------------------------------------------------------------------------
    def move_to_pending(self, job: TJob) -> None:
        job.status = "PENDING"
        self.pending_jobs.append(job)

    def remove_from_pending(self, job: TJob, event_time: int) -> None:
        job.status = "RUNNING"
        job.start_time = event_time
        job.end_time = event_time + job.duration
        job.pending_time = event_time - job.submit_time
        self.pending_jobs.remove(job)

    def update_pending_time(self, event_time: int) -> None:
        for job in self.pending_jobs:
            job.pending_time = int(event_time - job.submit_time)

    def move_to_runnable(self, job: TJob) -> None:
        job.status = "PENDING"
        job.start_time = MAXSIZE
        job.last_start_time = 0
        job.last_check_time = job.submit_time
        job.total_executed_time = 0
        job.executed_time = 0
        job.pending_time = 0
        job.last_pending_time = 0
        if self.multi_dlas:
            group = self.gpu_job.get(job.num_gpu)
            if group is None:
                group = self.gpu_job[job.num_gpu] = GpuJobGroup(job.num_gpu, self.num_queue, self.queue_limit)
            group.runnable_jobs.append(job)
            group.total_job += 1
        else:
            self.runnable_jobs.append(job)

    def end_job(self, job: TJob) -> None:
        """multi-dlas-gpu completion: leave the group's lists, free its GPUs."""
        group = self.gpu_job[job.num_gpu]
        group.runnable_jobs.remove(job)
        group.queues[job.q_id].remove(job)
        group.free_gpu = min(group.total_gpu, group.free_gpu + job.num_gpu)
        group.end_job += 1
        job.status = "END"

    def reserve_gpus(self, total_num: int) -> None:
        """Split the cluster's GPUs over the GPU-count groups in proportion to their demand (in
        whole jobs), never below what a group currently uses; left-overs go to the groups with
        unmet demand, smallest job size first."""
        groups = [g for g in self.gpu_job.values()]
        if not groups:
            return
        demand = {g.num_gpu: g.get_gpu_demands() for g in groups}
        total_demand = sum(demand.values())
        if total_demand == 0:
            return
        plan = {}
        for g in groups:
            share = demand[g.num_gpu] / total_demand * total_num
            want = min(demand[g.num_gpu], int(math.floor(share / g.num_gpu) * g.num_gpu))
            used = g.total_gpu - g.free_gpu
            plan[g.num_gpu] = max(want, used)
        spare = total_num - sum(plan.values())
        for g in sorted(groups, key=lambda x: x.num_gpu):
            while spare >= g.num_gpu and plan[g.num_gpu] + g.num_gpu <= demand[g.num_gpu]:
                plan[g.num_gpu] += g.num_gpu
                spare -= g.num_gpu
        for g in groups:
            g.set_reservation(plan[g.num_gpu])

    # -- gpu-demands mode --------------------------------------------------------------------------------
    def add_gpu_job(self, job: TJob) -> None:
        self.gpu_job[job.num_gpu] = self.gpu_job.get(job.num_gpu, 0) + 1

    def delete_gpu_job(self, job: TJob) -> None:
        self.gpu_job[job.num_gpu] = self.gpu_job.get(job.num_gpu, 0) - 1

    # -- placements recorded on a job ------------------------------------------------------------------------
    @staticmethod
    def create_single_node_placement(job: TJob, switch_id: int, node_id: int, num_gpu: int, num_cpu: int,
                                     mem: float = 0) -> float:
        job.placements.append({"switch": switch_id, "nodes": [
            {"id": node_id, "num_gpu": num_gpu, "num_cpu": num_cpu, "mem": mem, "network": 0, "tasks": []}]})
        return 0

    @staticmethod
    def create_multi_nodes_placement(job: TJob, switch_id: int, node_list: List[dict]) -> None:
        job.placements.append({"switch": switch_id, "nodes": node_list})

    def completion_check(self) -> bool:
        left = [j for j in self.job_list if j.status != "END"]
        return not left

    def total_jobs(self) -> int:
        return len(self.job_list)
