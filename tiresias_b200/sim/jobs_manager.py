"""The "application master" of the tick simulator (reference: core/jobs/jobs_manager.py:9-252, C10).

Feeds trace arrivals into the queues, starts / steps / preempts jobs and reports pending-time
statistics. Parity behaviour kept: arrival batches go to the *front* of list queues
(Q1); horus+ re-clusters every queued job on every ``insert`` without explicit positions, including
empty arrival batches (Q11); durations are ``minutes * scale_factor`` (Q4); a job completes when
its furthest task has processed ``>= duration`` ticks (Q6).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np

from . import util
from .infra import FACTOR
from .job import Job, order_for_schedule
from .kmeans import clusterize
from .trace import JobGenerator, JobTraceReader

LIVE_SCHEDULES = ("fifo", "horus", "horus+", "gandiva")


class JobsManager:
    def __init__(self, flags, job_queue_manager, rng=None, trace_reader: Optional[JobTraceReader] = None):
        self.job_queue_manager = job_queue_manager
        self.flags = flags
        self.rng = rng
        if trace_reader is not None:
            self.job_generator = trace_reader
            self.replay_trace = True
        elif flags.trace_file:
            self.job_generator = JobTraceReader(flags.trace_file)
            self.job_generator.prepare_jobs()
            self.replay_trace = True
        else:
            self.replay_trace = False
            self.job_generator = JobGenerator()
        self.running_jobs: Dict[str, Job] = {}
        self.finished_jobs: Dict[str, Job] = {}
        self.all_jobs: Dict[str, Job] = {}
        self.busy_nodes: List[str] = []
        self.interference_factor = FACTOR
        self.order = order_for_schedule(flags.schedule)
        self.apply_interference = getattr(flags, "mode", "parity") == "calibrated"

    # -- queue access ------------------------------------------------------------------------------
    def sort_job_trace(self) -> None:
        qs = self.job_queue_manager.queues
        for i, q in enumerate(qs):
            qs[i] = sorted(q, key=lambda x: int(x.submit_time))

    def get_next_job(self, delta, queue_idx: int = 0, job_in_queue: int = 0):
        j = self.job_queue_manager.get_next_job(queue_idx, job_in_queue)
        return j if j.submit_time <= delta else None

    def pop(self, delta, queue_idx: int = 0, job_in_queue: int = 0):
        jq = self.job_queue_manager
        if jq.length_of_queue(queue_idx) == 0:
            return None
        j = jq.pop(queue_idx, job_in_queue)
        if j.submit_time <= delta:
            return j
        jq.insert(j, queue_idx, job_in_queue)      # not yet submitted: put it back
        return None

    def get_insert_position(self, jobs) -> List[int]:
        if self.flags.schedule not in LIVE_SCHEDULES:
            raise NotImplementedError(self.flags.schedule)
        return list(range(len(jobs)))

    def get_queue_position(self, jobs) -> List[int]:
        sched = self.flags.schedule
        if sched in ("fifo", "gandiva", "horus"):
            return [0] * len(jobs)
        if sched == "horus+":
            if not jobs:
                return []
            _, queue_pos, loss = clusterize(jobs, k=len(self.job_queue_manager.queues), rng=self.rng)
            return queue_pos
        raise NotImplementedError(sched)

    def insert(self, jobs, queue_insert_position=None) -> None:
        jq = self.job_queue_manager
        recluster = self.flags.schedule == "horus+" and queue_insert_position is None
        if recluster:
            jobs = jq.pop_all_queuing_jobs() + jobs
        positions = self.get_insert_position(jobs)
        if queue_insert_position is None:
            queue_insert_position = self.get_queue_position(jobs)
        assert len(jobs) == len(queue_insert_position), (
            "Expected %d - Got %d" % (len(jobs), len(queue_insert_position)))
        for job, q_index, j_index in zip(jobs, queue_insert_position, positions):
            jq.insert(job, q_index, j_index)

    # -- statistics ----------------------------------------------------------------------------------
    def add_pending_time(self) -> None:
        for q in self.job_queue_manager.queues:
            for j in q:
                j.step_pending()

    def pending_time_infos(self):
        """(mean, median, max) pending time over queued jobs; the mean's denominator carries the
        reference's ``+1e-9`` and the median of nothing is NaN."""
        total, count, longest = 0, 0, 0
        every = []
        for q in self.job_queue_manager.queues:
            for j in q:
                p = j.pending_time
                longest = max(p, longest)
                every.append(p)
                total += p
            count += len(q)
        if every:
            median = float(np.median(every))
        else:
            median = float("nan")
        return total / (count + 1e-9), median, longest

    # -- lifecycle -------------------------------------------------------------------------------------
    def step(self) -> None:
        self.add_pending_time()
        for j in self.running_jobs.values():
            j.step()
        if self.flags.schedule == "horus+":
            self.job_queue_manager.update_credits()

    def preempt(self, job_id, infrastructure) -> None:
        job = self.running_jobs.pop(job_id)
        seen = set()
        for t, n in job.tasks_running_on.items():
            node = infrastructure.nodes[n]
            if n not in seen:
                node.placed_jobs.pop(job_id)
                seen.add(n)
            task = node.running_tasks.pop(t, None)
            if task is None:
                continue
            relieved = node.release_allocated_resources(task, reserved=True)
            if relieved:
                if t in relieved and task.interfered:
                    task.interfered = False
                    task.duration = task.duration * (1 - self.interference_factor)
                    relieved.pop(t)
                self.reset_interference(relieved)
        job.preempted()
        self.insert([job])

    def reset_interference(self, relieved: Dict[str, str]) -> None:
        for t, j in relieved.items():
            running = self.running_jobs.get(j)
            if running is None:
                continue
            task = running.tasks[t]
            if task.interfered:
                task.interfered = False
                task.duration = task.duration * (1 - self.interference_factor)

    def start_job(self, node, job_id, delta_time) -> bool:
        job, started = node.execute_job(job_id, delta_time, apply_interference=self.apply_interference)
        assert started > 0
        if node.node_id not in self.busy_nodes:
            self.busy_nodes.append(node.node_id)
        if job is None:
            return False
        if job.job_id not in self.running_jobs:
            self.running_jobs[job.job_id] = job
        return True

    # -- counters ----------------------------------------------------------------------------------------
    def remaining_jobs(self, delta_time=None) -> int:
        if self.replay_trace:
            return self.job_generator.remaining_jobs()
        raise NotImplementedError()

    def queuing_jobs(self, delta_time=None) -> int:
        return self.job_queue_manager.total_jobs(delta_time)

    def total_jobs(self, delta_time) -> int:
        queued = self.job_queue_manager.total_jobs(delta_time)
        return self.remaining_jobs() + queued if self.replay_trace else queued

    def total_finished_jobs(self) -> int:
        return len(self.finished_jobs)

    def set_measured_profiles(self, calibration) -> None:
        """K7, calibrated mode: jobs whose trace row names a model that ``tiresias_b200.calib`` measured
        take that model's GPU utilisation (NVML) and peak device memory instead of the trace's columns,
        and their devices report the measured utilisation instead of drawing one (``Device.measured``)."""
        cal = calibration or {}
        self.measured_util = dict(cal.get("gpu_util") or {})
        self.measured_mem = dict(cal.get("peak_mem_bytes") or {})

    def _profile(self, row):
        """(util_avg, util_max, mem_max_MiB, mem_avg_MiB, measured?) of a trace row."""
        u_avg, u_max = row.gpu_utilization_avg, row.gpu_utilization_max
        m_max = util.convert_bytes(row.memory_max, unit="MiB")
        m_avg = util.convert_bytes(row.memory_avg, unit="MiB")
        measured = False
        name = getattr(row, "model_name", None)
        if self.apply_interference and name:
            mu = getattr(self, "measured_util", {}).get(name)
            if mu:
                u_avg, u_max, measured = float(mu["avg"]), float(mu.get("max", mu["avg"])), True
            mm = getattr(self, "measured_mem", {}).get(name)
            if mm:
                m_max = m_avg = util.convert_bytes(mm, unit="MiB")
                measured = True
        return u_avg, u_max, m_max, m_avg, measured

    def gen_jobs(self, delta_time, scale_factor=1) -> int:
        rows = self.job_generator.generate_jobs(delta_time)
        jobs = []
        for row in rows:
            u_avg, u_max, m_max, m_avg, measured = self._profile(row)
            j = Job(row.index, row.minutes * scale_factor, row.normalized_time, row.gpu_per_container,
                    gpu_utilization_avg=u_avg,
                    gpu_utilization_max=u_max,
                    gpu_memory_max=m_max,
                    gpu_memory_avg=m_avg,
                    total_gpus=row.used_gpus, order=self.order, model_name=row.model_name,
                    iterations=row.iterations)
            j.arrival_time = row.normalized_time
            if measured:
                for t in j.tasks.values():
                    t.measured_profile = True
            jobs.append(j)
            self.all_jobs[j.job_id] = j
        self.insert(jobs)
        return len(rows)

    def prepare_finish_tasks(self, current_time) -> List[Job]:
        """Jobs whose furthest task has processed ``duration`` ticks. Parity mode compares against
        ``Job.duration`` only (as the reference does, which is one reason its interference stretch is
        inert); calibrated mode honours the per-task durations that co-location has stretched."""
        if not self.apply_interference:
            return [j for j in self.running_jobs.values() if not j.time_processed() < j.duration]
        out = []
        for j in self.running_jobs.values():
            need = max([j.duration] + [t.duration for t in j.tasks.values()])
            if not j.time_processed() < need:
                out.append(j)
        return out

    def set_interference_factor(self, factor: float, infrastructure=None) -> None:
        """Replace the reference's constant 0.2 (infra/interference.py:1) by a measured slowdown."""
        self.interference_factor = factor
        if infrastructure is not None:
            for node in infrastructure.nodes.values():
                node.interference_factor = factor
