"""Job / task state of the tick-based simulator (reference: core/jobs/job.py:6-201, C9; ordering
mixins core/jobs/base_factory.py:2-30, C10b).

A job asks for ``total_gpus`` GPUs in workers of ``gpu_p_worker`` GPUs each; every worker is one
:class:`Task` (12 CPUs, 60 memory units, named ``<job_id>_worker<i>``). Progress lives in the
tasks (``time_processed``), so a preempted job resumes where it stopped.

Instead of choosing a base class at import time (the reference's ``BASE_OBJ`` trick, which allows
one ordering per process), each job carries an ``order`` tag that ``__lt__`` dispatches on:
``"utilization"`` (horus family: smaller average GPU utilisation first, a falsy utilisation never
compares smaller — SURVEY Q10) or ``"pending"`` (longer pending first).
"""
from __future__ import annotations

from typing import Dict, List, Tuple

ORDER_UTILIZATION = "utilization"
ORDER_PENDING = "pending"

CPUS_PER_TASK = 12      # core/jobs/job.py:103
MEMORY_PER_TASK = 60    # core/jobs/job.py:104


def order_for_schedule(schedule: str) -> str:
    """horus* schedules order their heaps by utilisation, everything else by pending time."""
    return ORDER_UTILIZATION if schedule.startswith("horus") else ORDER_PENDING


class _Ordered:
    __slots__ = ()

    def __lt__(self, other) -> bool:
        if self.order == ORDER_UTILIZATION:
            if self.gpu_utilization_avg:
                return self.gpu_utilization_avg < other.gpu_utilization_avg
            return False
        return self.pending_time > other.pending_time


class Task(_Ordered):
    """One worker of a job."""

    __slots__ = ("job_id", "task_id", "cpu", "mem", "gpu", "start_time", "gpu_utilization_avg",
                 "gpu_utilization_max", "gpu_memory_avg", "gpu_memory_max", "pending_time", "duration",
                 "started", "running", "finished", "interfered", "migration_count", "time_processed",
                 "order", "is_ps", "measured_profile")

    def __init__(self, job_id, task_id, duration, gpu_utilization_avg=0, gpu_utilization_max=0,
                 gpu_mem_avg=0, gpu_mem_max=0, cpu=0, mem=0, gpu=0, order=ORDER_PENDING):
        self.job_id = str(job_id)
        self.task_id = str(task_id)
        self.cpu, self.mem, self.gpu = cpu, mem, gpu
        self.start_time = 0
        self.gpu_utilization_avg = gpu_utilization_avg
        self.gpu_utilization_max = gpu_utilization_max
        self.gpu_memory_avg = gpu_mem_avg
        self.gpu_memory_max = gpu_mem_max
        self.pending_time = 0
        self.duration = duration
        self.started = self.running = self.finished = self.interfered = False
        self.migration_count = 0
        self.time_processed = 0
        self.order = order
        self.is_ps = False
        self.measured_profile = False      # calibrated mode: utilisation/memory measured on B200 (K7)

    def execute(self, delta_time) -> None:
        self.start_time = delta_time
        self.migration_count += 1
        self.started = self.running = True

    def preempted(self) -> None:
        self.running = False
        self.pending_time = 0

    def step(self) -> None:
        if self.running:
            self.time_processed += 1

    def __repr__(self) -> str:
        return f"Task({self.task_id}, gpu={self.gpu}, done={self.time_processed}/{self.duration})"


class Job(_Ordered):
    """A training job (synchronous SGD; README.md:3-7 assumptions)."""

    def __init__(self, job_id, duration, submit_time, gpu_p_worker=0, gpu_utilization_avg=0,
                 gpu_utilization_max=0, gpu_memory_max=0, gpu_memory_avg=0, total_gpus=0,
                 order=ORDER_PENDING, model_name=None, iterations=None, model_size=None):
        self.job_id = str(job_id)
        self.started = self.running = self.finished = False
        self.failed_schedule = 0
        self.start_time = 0
        self.duration = duration
        self.submit_time = int(submit_time)
        self.pending_time = 0.0
        self.migration_count = 0
        self.worker_count = total_gpus // gpu_p_worker
        self.gpu_per_worker = gpu_p_worker
        self.gpus = total_gpus
        self.task_count = int(self.worker_count)
        self.task_id = ["worker" + str(i) for i in range(self.task_count)]
        self.gpu_mem_avg = gpu_memory_avg
        self.gpu_mem_max = gpu_memory_max
        self.gpu_utilization_avg = gpu_utilization_avg
        self.gpu_utilization_max = gpu_utilization_max
        self.cpus_per_task = CPUS_PER_TASK
        self.memory_per_task = MEMORY_PER_TASK
        self.tasks_running_on: Dict[str, str] = {}
        self.tasks_finished = 0
        self.order = order
        # execution-backend / network-model metadata (absent from the reference's live Job, which
        # is why its --enable_network_costs path raises)
        self.model_name = model_name
        self.iterations = iterations
        self.model_size = model_size
        self.ps_count = 0
        # per-job record for jobs.csv (the reference's live path never logs per-job rows)
        self.first_start_time = None
        self.end_time = None
        self.preempt_count = 0
        self.promote_count = 0
        self.total_pending = 0.0
        self.tasks: Dict[str, Task] = self.setup_tasks()

    # -- sizing ----------------------------------------------------------------------------------
    def total_cpus_required(self) -> int:
        return self.cpus_per_task * self.task_count

    def total_mem_required(self) -> int:
        return self.memory_per_task * self.task_count

    def setup_tasks(self) -> Dict[str, Task]:
        tasks = {}
        for name in self.task_id:
            t = Task(self.job_id, self.job_id + "_" + name, self.duration,
                     gpu_utilization_avg=self.gpu_utilization_avg,
                     gpu_utilization_max=self.gpu_utilization_max, gpu_mem_avg=self.gpu_mem_avg,
                     gpu_mem_max=self.gpu_mem_max, cpu=self.cpus_per_task, mem=self.memory_per_task,
                     gpu=self.gpu_per_worker, order=self.order)
            tasks[t.task_id] = t
        return tasks

    # -- state ------------------------------------------------------------------------------------
    def is_waiting(self) -> bool:
        return not self.running

    def is_preempted(self) -> bool:
        return self.started and not self.running

    def task_finished(self, t_id: str) -> None:
        t = self.tasks.get(t_id)
        if t is not None and not t.finished:
            t.finished = True
            self.tasks_finished += 1

    def try_finished(self) -> bool:
        if self.finished:
            return True
        if self.tasks_finished == self.task_count:
            self.running = False
            self.finished = True
            return True
        return False

    def step_pending(self) -> None:
        self.pending_time += 1
        self.total_pending += 1
        for t in self.tasks.values():
            t.pending_time += 1

    def try_execute(self, delta_time) -> Tuple[bool, int]:
        """The job runs only once every one of its tasks runs."""
        live = sum(1 for t in self.tasks.values() if t.running and not t.finished)
        if live == self.task_count:
            self.start_time = delta_time
            if self.first_start_time is None:
                self.first_start_time = delta_time
            self.migration_count += 1
            self.started = self.running = True
            return True, live
        return False, live

    def preempted(self) -> None:
        self.running = False
        for t in self.tasks.values():
            t.preempted()
        self.pending_time = 0
        self.preempt_count += 1

    def step(self) -> None:
        if self.running:
            for t in self.tasks.values():
                t.step()

    def time_processed(self):
        done = 0
        for t in self.tasks.values():
            if t.time_processed > done:
                done = t.time_processed
        return done

    def add_network_costs(self, extra_s) -> None:
        self.duration += extra_s

    def is_distributed(self) -> bool:
        """More than one worker. (The reference tests a ``ps_count`` attribute that its Job never
        defines, core/jobs/job.py:196-197, so its network-cost path raises; here every GPU hosts a
        worker and a PS shard, infra/switch.py:100-106, hence workers > 1 <=> PS shards > 1.)"""
        return self.task_count > 1

    def restart(self) -> None:
        self.running = True
        self.migration_count += 1

    def __repr__(self) -> str:
        return f"Job({self.job_id}, gpus={self.gpus}, dur={self.duration}, submit={self.submit_time})"


def task_ids(job: Job) -> List[str]:
    return list(job.tasks.keys())
