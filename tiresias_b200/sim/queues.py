"""Waiting queues (reference: core/jobs/job_queue_manager.py:9-154, C10a).

``num_queue`` queues; each is a binary heap (ordered by the jobs' ``__lt__``) when the schedule is
of the horus family, otherwise a plain list. ``queue_limit`` here is a *capacity* (9999 per queue),
not an MLFQ threshold — the MLFQ thresholds of the Tiresias schedulers live in
:mod:`tiresias_b200.sim.legacy`. Credits drive horus+'s choice of queue: ``median(pending) * len``
(``len`` while the median is below 1), plus one on every insert; the highest credit wins, lowest
index on ties.
"""
from __future__ import annotations

from heapq import heappop, heappush
from typing import List, Optional

import numpy as np

QUEUE_CAPACITY = 9999


class JobQueueManager:
    def __init__(self, flags, file_path: Optional[str] = None):
        self.flags = flags
        self.file_path = file_path
        self.num_queue = flags.num_queue
        self.queues: List[list] = [[] for _ in range(self.num_queue)]
        self.queue_limit = [QUEUE_CAPACITY] * self.num_queue
        self.queue_credits: List[float] = [0] * self.num_queue
        self.queues_is_pq = [flags.schedule.startswith("horus")] * self.num_queue

    # -- README-schema loader (the reference's version is dead code that cannot run; D15) ----------
    def parse_job_file(self) -> List[dict]:
        from .trace import read_tiresias_trace
        return read_tiresias_trace(self.file_path)

    # -- sizes -------------------------------------------------------------------------------------
    def length_of_queue(self, qidx: int) -> int:
        return len(self.queues[qidx])

    def total_jobs(self, delta_time=-1) -> int:
        return sum(len(q) for q in self.queues)

    def _can_add(self, queue_idx: int) -> bool:
        return len(self.queues[queue_idx]) < self.queue_limit[queue_idx]

    def _add_to_job_queue(self, new_job, queue_idx: Optional[int] = None) -> None:
        q = 0 if queue_idx is None else queue_idx
        if not self._can_add(q):
            raise ArithmeticError(f"queue {q} is full")
        if self.queues_is_pq[q]:
            heappush(self.queues[q], new_job)
        else:
            self.queues[q].append(new_job)

    # -- access ------------------------------------------------------------------------------------
    def get_next_job(self, queue_idx: int = 0, job_in_queue: int = 0):
        return self.queues[queue_idx][job_in_queue]

    def pop(self, queue_idx: int = 0, job_in_queue: int = 0):
        q = self.queues[queue_idx]
        if self.queues_is_pq[queue_idx]:
            if not q:
                raise ValueError("Length of queue %d is %d" % (queue_idx, len(q)))
            return heappop(q)
        return q.pop(job_in_queue)

    def pop_all_queuing_jobs(self) -> list:
        """Empties every queue (heap order within a queue, queues in index order)."""
        jobs = []
        for q_idx in range(self.num_queue):
            for _ in range(len(self.queues[q_idx])):
                jobs.append(self.pop(q_idx))
        return jobs

    def insert(self, job, queue_idx: int = 0, job_in_queue: int = 0) -> None:
        self.queue_credits[queue_idx] = self.queue_credits[queue_idx] + 1
        if self.queues_is_pq[queue_idx]:
            heappush(self.queues[queue_idx], job)
        else:
            self.queues[queue_idx].insert(job_in_queue, job)

    # -- credits -----------------------------------------------------------------------------------
    def queue_idx_by_credit(self) -> int:
        return int(np.argmax(self.queue_credits))

    def update_credits(self) -> None:
        for q in range(self.num_queue):
            jobs = self.queues[q]
            if jobs:
                med = max(np.median([j.pending_time for j in jobs]), 0)
                self.queue_credits[q] = len(jobs) if med < 1 else med * len(jobs)
            else:
                self.queue_credits[q] = 0
