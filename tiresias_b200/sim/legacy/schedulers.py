"""The event-driven Tiresias schedulers (reference: dead code run_sim.py:98-1647, D1-D10; it cannot
run as shipped because its ``JOBS`` / ``CLUSTER`` / ``LOG`` globals are never created — this is a
restatement checked against hand-computed micro-traces, tests/test_sim_legacy.py).

Families:

* non-preemptive event loop — ``fifo`` / ``fjf`` (:func:`sim_job_events`), ``lpjf``
  (:func:`longest_pending_first`): completions free resources, pending jobs are tried in order
  until the first failure, arrivals start or pend.
* preemptive re-placement — ``sjf`` (smallest ``num_gpu`` first), ``shortest``, ``shortest-gpu``,
  ``shortest-expected`` (:func:`priority_replace`): at every arrival/completion all runnable jobs are
  sorted by the policy key, the cluster is emptied and re-filled in that order; running jobs that no
  longer fit are preempted, pending ones that now fit resume.
* MLFQ / discretised LAS — ``dlas``, ``dlas-gpu``, ``dlas-gpu-gittins``, ``dlas-gpu-pack``
  (:func:`dlas`), ``multi-dlas-gpu`` (:func:`multi_dlas`): jobs enter Q0, are demoted one level when
  their attained service (seconds, or GPU-seconds with ``-gpu``) reaches ``queue_limit[q]``, may be
  promoted back to Q0 when starved; GPUs are handed out by *counting* in queue order; the loop also
  wakes at the next demotion / promotion instant.
* ``gittins`` (:func:`gittins`): one logical queue ordered by the Gittins index of attained service,
  re-evaluated at least every ``gittins_delta`` units.
* ``gandiva-legacy`` (:func:`gandiva_legacy`): 10-second stepping with node sets and time slicing.
* ``gpu-demands`` (:func:`gpu_demands`): no placement; logs demanded GPUs per job size.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, List, Optional, Tuple



class _LazyTorch:
    """``torch`` on first use: the priority math (LAS queues, Gittins ranks) is in PyTorch, but importing it
    costs seconds that fifo / sjf / shortest runs never need."""

    def __getattr__(self, name):
        import torch as _torch
        globals()["torch"] = _torch
        return getattr(_torch, name)


torch = _LazyTorch()

from . import gittins as gi
from .jobs import MAXSIZE, JobRegistry, TJob

NOT_LARGE_ENOUGH = "This cluster is not large enough to run the job"


class Sim:
    """Bundle handed to every scheduler: flags + the JOBS / CLUSTER / LOG triple."""

    def __init__(self, flags, jobs: JobRegistry, cluster, log, print_fn=None):
        self.flags, self.jobs, self.cluster, self.log = flags, jobs, cluster, log
        self.print_fn = print_fn or (lambda *_: None)
        self.events = 0
        self.last_time = 0

    def mark(self, event_time: int) -> None:
        self.events += 1
        self.last_time = event_time


# ---- shared pieces -----------------------------------------------------------------------------------

def _next_event(jobs: JobRegistry, end_events: List[dict], pop_start: bool) -> Tuple[int, dict]:
    """Earliest of the next arrival and the next completion. When both fall on the same instant the
    arrival event also carries the completions. ``pop_start`` removes the chosen arrival event
    right away (the sjf/shortest loops); the MLFQ loops pop it after processing instead."""
    start_time = jobs.job_events[0]["time"] if jobs.job_events else MAXSIZE
    end_time = end_events[0]["time"] if end_events else MAXSIZE
    if end_time < start_time:
        return end_time, end_events[0]
    if end_time > start_time:
        return start_time, (jobs.job_events.pop(0) if pop_start else jobs.job_events[0])
    if end_time != MAXSIZE:
        event = jobs.job_events.pop(0) if pop_start else jobs.job_events[0]
        event["end_jobs"] = end_events[0]["end_jobs"]
        return start_time, event
    return MAXSIZE, {"time": MAXSIZE}


def _earliest_completion(runnable: List[TJob], event_time: int) -> Optional[dict]:
    """The single next completion event among running jobs (ties share it)."""
    best, event = MAXSIZE, None
    for j in runnable:
        if j.status != "RUNNING":
            continue
        end = int(event_time + (j.duration - j.total_executed_time))
        if end < best:
            best, event = end, {"time": end, "end_jobs": [j]}
        elif end == best:
            event["end_jobs"].append(j)
    return event


RESUME_OVERHEAD = 0      # time units added to a job each time it resumes after a preemption (0 = reference)


def _apply_switches(preempt_jobs: List[TJob], run_jobs: List[TJob], event_time: Optional[int]) -> None:
    for j in preempt_jobs:
        j.status = "PENDING"
        j.preempt = int(j.preempt + 1)
    for j in run_jobs:
        j.status = "RUNNING"
        j.resume = int(j.resume + 1)
        if RESUME_OVERHEAD and j.preempt > 0 and j.total_executed_time > 0:
            # The reference prices preemption at zero (progress simply continues). On real
            # GPUs a resumed job rebuilds its model, re-exchanges IPC handles and reloads its PS shards:
            # the measured restart cost (calibration "resume_overhead") is charged as extra work.
            j.duration = int(j.duration + RESUME_OVERHEAD)
            j.remaining_time = int(j.duration - j.total_executed_time)
        if event_time is not None and j.start_time == MAXSIZE:
            j.start_time = event_time


def _pending_to_back(queues: List[List[TJob]]) -> None:
    """Within every queue keep running jobs in front, pending ones (in order) behind them."""
    for q in queues:
        waiting = [j for j in q if j.status == "PENDING"]
        for j in waiting:
            q.remove(j)
        q.extend(waiting)


def _attained(job: TJob, gputime: bool) -> int:
    return int(job.executed_time * job.num_gpu) if gputime else int(job.executed_time)


# ---- non-preemptive ------------------------------------------------------------------------------------

def _start_now(sim: Sim, job: TJob, event_time: int) -> None:
    job.status = "RUNNING"
    job.start_time = event_time
    job.end_time = event_time + job.duration
    job.pending_time = event_time - job.submit_time
    sim.jobs.add_job_end_event(job)


def sim_job_events(sim: Sim) -> None:
    """fifo / fjf (run_sim.py:1533-1599). Pending jobs are kept sorted by ``num_gpu`` after every
    event and tried front to back, stopping at the first that does not fit."""
    jobs, cluster, log = sim.jobs, sim.cluster, sim.log
    while len(jobs.job_events) + len(jobs.pending_jobs) > 0:
        if not jobs.job_events:
            sim.print_fn(NOT_LARGE_ENOUGH)
            break
        event = jobs.job_events[0]
        now = event["time"]
        for j in event["end_jobs"]:
            cluster.release_job_res(j)
            log.job_complete(j, now)
        for j in jobs.pending_jobs:                 # removal while iterating skips the next entry,
            if cluster.try_get_job_res(j):          # exactly as the reference's loop does
                jobs.remove_from_pending(j, now)
                jobs.add_job_end_event(j)
            else:
                break
        for j in event["start_jobs"]:
            if cluster.try_get_job_res(j):
                _start_now(sim, j, now)
            else:
                jobs.move_to_pending(j)
        jobs.pending_jobs.sort(key=lambda j: j.num_gpu)
        jobs.job_events.remove(event)
        jobs.job_events.sort(key=lambda e: e["time"])
        log.checkpoint(jobs, now)
        sim.mark(now)


def longest_pending_first(sim: Sim) -> None:
    """lpjf (run_sim.py:1474-1531): arrivals always queue; longest-waiting first until one fails."""
    jobs, cluster, log = sim.jobs, sim.cluster, sim.log
    while len(jobs.job_events) + len(jobs.pending_jobs) > 0:
        if not jobs.job_events:
            sim.print_fn(NOT_LARGE_ENOUGH)
            break
        event = jobs.job_events[0]
        now = event["time"]
        for j in event["end_jobs"]:
            cluster.release_job_res(j)
            log.job_complete(j, now)
        for j in event["start_jobs"]:
            jobs.move_to_pending(j)
        jobs.update_pending_time(now)
        jobs.pending_jobs.sort(key=lambda j: j.pending_time, reverse=True)
        for j in jobs.pending_jobs:
            if cluster.try_get_job_res(j):
                jobs.remove_from_pending(j, now)
                jobs.add_job_end_event(j)
            else:
                break
        jobs.job_events.remove(event)
        jobs.job_events.sort(key=lambda e: e["time"])
        log.checkpoint(jobs, now)
        sim.mark(now)


# ---- preemptive re-placement -------------------------------------------------------------------------------

def priority_replace(sim: Sim, policy: str) -> None:
    """sjf / shortest / shortest-gpu / shortest-expected (run_sim.py:161-430)."""
    jobs, cluster, log = sim.jobs, sim.cluster, sim.log
    gputime = policy == "shortest-gpu"
    keys: Dict[str, Callable[[TJob], float]] = {
        "sjf": lambda j: j.num_gpu,
        "shortest": lambda j: j.remaining_time,
        "shortest-gpu": lambda j: j.remaining_gputime,
        "shortest-expected": lambda j: j.remaining_expected,
    }
    key = keys[policy]
    end_events: List[dict] = []
    while len(jobs.job_events) + len(jobs.runnable_jobs) > 0:
        if len(jobs.job_events) + len(end_events) == 0:
            sim.print_fn(NOT_LARGE_ENOUGH)
            break
        now, event = _next_event(jobs, end_events, pop_start=True)
        for j in event.get("end_jobs", ()):
            cluster.release_job_res(j)
            log.job_complete(j, now)
            jobs.runnable_jobs.remove(j)
        for j in event.get("start_jobs", ()):
            jobs.move_to_runnable(j)
            if policy == "shortest-expected":
                j.remaining_expected = gi.cal_shortest_expected_remaining(jobs.job_dist_data, 0)
            j.remaining_time = j.duration
            j.remaining_gputime = int(j.remaining_time * j.num_gpu)
        for j in list(jobs.runnable_jobs):
            dt = int(now - j.last_check_time)
            if j.status == "RUNNING":
                j.total_executed_time = int(j.total_executed_time + dt)
                j.last_check_time = now
                j.remaining_time = int(j.duration - j.total_executed_time)
                if policy == "shortest-expected":
                    j.remaining_expected = gi.cal_shortest_expected_remaining(
                        jobs.job_dist_data, j.total_executed_time)
                if gputime:
                    j.remaining_gputime = int(j.remaining_time * j.num_gpu)
            elif j.status == "PENDING":
                j.pending_time = int(j.pending_time + dt)
                j.last_check_time = now
        jobs.runnable_jobs.sort(key=key)
        run_jobs, preempt_jobs = [], []
        cluster.empty_infra()
        for j in jobs.runnable_jobs:
            if j.status == "RUNNING":
                del j.placements[:]
            if cluster.try_get_job_res(j):
                if j.start_time == MAXSIZE:
                    j.start_time = now
                if j.status == "PENDING":
                    run_jobs.append(j)
            elif j.status == "RUNNING":
                preempt_jobs.append(j)
        _apply_switches(preempt_jobs, run_jobs, None)
        del end_events[:]
        by_time: Dict[int, dict] = {}
        for j in jobs.runnable_jobs:
            if j.status == "RUNNING":
                end = int(now + j.remaining_time)
                ev = by_time.get(end)
                if ev is None:
                    ev = by_time[end] = {"time": end, "end_jobs": []}
                    end_events.append(ev)
                ev["end_jobs"].append(j)
        end_events.sort(key=lambda e: e["time"])
        log.checkpoint(jobs, now)
        sim.mark(now)


# ---- MLFQ / LAS ----------------------------------------------------------------------------------------------

def _las_clock_update(job: TJob, now: int, gputime: bool, num_queue: int, limits: List[int],
                      queues: List[List[TJob]], solve_starvation: float) -> None:
    """Advance one job's clocks to ``now`` and apply demotion (running) or promotion (pending)."""
    dt = int(now - job.last_check_time)
    if job.status == "RUNNING":
        job.total_executed_time = int(job.total_executed_time + dt)
        job.executed_time = int(job.executed_time + dt)
        job.last_check_time = now
        q = job.q_id
        if q < num_queue - 1 and _attained(job, gputime) >= limits[q]:
            job.q_id = q + 1
            queues[q + 1].append(job)
            queues[q].remove(job)
    elif job.status == "PENDING":
        job.last_check_time = now
        job.pending_time = int(job.pending_time + dt)
        if job.executed_time > 0:
            job.last_pending_time = int(job.last_pending_time + dt)
        if (solve_starvation > 0 and job.q_id > 0 and job.total_executed_time > 0
                and job.executed_time > 0
                and job.last_pending_time >= int(job.executed_time * solve_starvation)):
            job.executed_time = 0
            job.last_pending_time = 0
            queues[0].append(job)
            queues[job.q_id].remove(job)
            job.q_id = 0
            job.promote = int(job.promote + 1)


def _next_jump(runnable: List[TJob], now: int, gputime: bool, num_queue: int, limits: List[int],
               solve_starvation: float) -> int:
    """Next instant at which some job would change queue. The GPU-time form divides the *remaining
    executed-seconds gap* by num_gpu (run_sim.py:929), which can overshoot the true instant; the
    demotion test itself is exact and runs at every event."""
    best = MAXSIZE
    for j in runnable:
        if j.status == "RUNNING":
            q = j.q_id
            if q < num_queue - 1:
                if gputime:
                    t = int(math.ceil((limits[q] - j.executed_time) / j.num_gpu) + now)
                else:
                    t = int(limits[q] - j.executed_time + now)
                best = min(best, t)
        elif j.status == "PENDING":
            if solve_starvation > 0 and j.q_id > 0 and j.total_executed_time and j.executed_time > 0:
                diff = int(j.executed_time * solve_starvation - j.last_pending_time)
                if diff > 0:
                    best = min(best, int(diff + now))
    return best


def dlas(sim: Sim, gputime: bool = False, solve_starvation: float = 0) -> None:
    """dlas / dlas-gpu / dlas-gpu-gittins / dlas-gpu-pack (run_sim.py:663-946, :1204-1472)."""
    jobs, cluster, log, flags = sim.jobs, sim.cluster, sim.log, sim.flags
    with_gittins = flags.schedule == "dlas-gpu-gittins"
    pack = flags.schedule == "dlas-gpu-pack"
    end_events: List[dict] = []
    next_job_jump = MAXSIZE
    while len(jobs.job_events) + len(jobs.runnable_jobs) > 0:
        if len(jobs.job_events) + len(end_events) == 0:
            sim.print_fn(NOT_LARGE_ENOUGH)
            break
        now, event = _next_event(jobs, end_events, pop_start=False)
        if now > next_job_jump:
            now, event = next_job_jump, {}
        for j in event.get("end_jobs", ()):
            cluster.release_job_res(j)
            log.job_complete(j, now)
            jobs.runnable_jobs.remove(j)
            jobs.queues[j.q_id].remove(j)
        if "start_jobs" in event:
            for j in event["start_jobs"]:
                jobs.move_to_runnable(j)
                j.q_id = 0
                jobs.queues[0].append(j)
            jobs.job_events.pop(0)
        for j in list(jobs.runnable_jobs):
            _las_clock_update(j, now, gputime, jobs.num_queue, jobs.queue_limit, jobs.queues, solve_starvation)
        if with_gittins and jobs.runnable_jobs:
            att = torch.tensor([_attained(j, gputime) for j in jobs.runnable_jobs], dtype=torch.int64)
            for j, r in zip(jobs.runnable_jobs, gi.gittins_ranks(jobs.job_dist_data, att).tolist()):
                j.rank = r
        cluster.empty_infra()
        run_jobs, preempt_jobs = [], []
        for q in jobs.queues:
            if with_gittins:
                q.sort(key=lambda j: (j.rank, j.r_submit_time), reverse=True)
            for j in q:
                if pack:
                    fits = cluster.free_gpu_util(j)
                    if fits:
                        del j.gpus[:]
                        cluster.dlas_pack_get_gpu_util(j)
                else:
                    fits = cluster.free_gpu >= j.num_gpu
                    if fits:
                        cluster.free_gpu = int(cluster.free_gpu - j.num_gpu)
                if fits:
                    if j.status == "PENDING":
                        run_jobs.append(j)
                elif j.status == "RUNNING":
                    preempt_jobs.append(j)
        _apply_switches(preempt_jobs, run_jobs, now)
        _pending_to_back(jobs.queues)
        del end_events[:]
        nxt = _earliest_completion(jobs.runnable_jobs, now)
        if nxt is not None:
            end_events.append(nxt)
        next_job_jump = _next_jump(jobs.runnable_jobs, now, gputime, jobs.num_queue, jobs.queue_limit,
                                   solve_starvation)
        log.checkpoint(jobs, now)
        sim.mark(now)


def multi_dlas(sim: Sim, gputime: bool = False, solve_starvation: float = 0,
               reserve_interval: int = 600) -> None:
    """multi-dlas-gpu (run_sim.py:432-661): one MLFQ per GPU-count class; the GPUs reserved for each
    class are re-balanced to demand every ``reserve_interval`` seconds."""
    jobs, cluster, log = sim.jobs, sim.cluster, sim.log
    jobs.multi_dlas = True
    end_events: List[dict] = []
    next_job_jump = MAXSIZE
    next_reserve = reserve_interval
    while len(jobs.job_events) + sum(len(g.runnable_jobs) for g in jobs.gpu_job.values()) > 0:
        if len(jobs.job_events) + len(end_events) == 0:
            sim.print_fn(NOT_LARGE_ENOUGH)
            break
        now, event = _next_event(jobs, end_events, pop_start=False)
        if now > next_job_jump:
            now, event = next_job_jump, {}
        if now > next_reserve:
            now, event = next_reserve, {}
            next_reserve += reserve_interval
            jobs.reserve_gpus(cluster.num_gpu)
        for j in event.get("end_jobs", ()):
            jobs.end_job(j)
            log.job_complete(j, now)
        if "start_jobs" in event:
            fresh_group = False
            for j in event["start_jobs"]:
                known = j.num_gpu in jobs.gpu_job
                jobs.move_to_runnable(j)
                j.q_id = 0
                jobs.gpu_job[j.num_gpu].queues[0].append(j)
                fresh_group = fresh_group or not known
            jobs.job_events.pop(0)
            if fresh_group:                        # a new class must get GPUs before its first pass
                jobs.reserve_gpus(cluster.num_gpu)
        del end_events[:]
        best_end, end_event = MAXSIZE, None
        next_job_jump = MAXSIZE
        for group in jobs.gpu_job.values():
            if not group.runnable_jobs:
                continue
            for j in list(group.runnable_jobs):
                _las_clock_update(j, now, gputime, group.num_queue, group.queue_limit, group.queues,
                                  solve_starvation)
            group.empty_gpu_alloc()
            run_jobs, preempt_jobs = [], []
            for q in group.queues:
                for j in q:
                    if group.alloc_free_gpus(j.num_gpu):
                        if j.status == "PENDING":
                            run_jobs.append(j)
                    elif j.status == "RUNNING":
                        preempt_jobs.append(j)
            _apply_switches(preempt_jobs, run_jobs, now)
            _pending_to_back(group.queues)
            cand = _earliest_completion(group.runnable_jobs, now)
            if cand is not None:
                if cand["time"] < best_end:
                    best_end, end_event = cand["time"], cand
                elif cand["time"] == best_end:
                    end_event["end_jobs"].extend(cand["end_jobs"])
            next_job_jump = min(next_job_jump, _next_jump(group.runnable_jobs, now, gputime, group.num_queue,
                                                          group.queue_limit, solve_starvation))
        if end_event is not None:
            end_events.append(end_event)
        log.checkpoint_multi_dlas_gpu(jobs, now)
        sim.mark(now)
    jobs.completion_check()


def gittins(sim: Sim, gputime: bool = False, descending: bool = False, clean: bool = False) -> None:
    """gittins (run_sim.py:956-1202). ``descending=False`` is the reference's sort direction
    (ascending rank, run_sim.py:1082); its own dlas-gpu-gittins variant sorts descending (:809) —
    the ambiguity is kept selectable (``--gittins_desc``). ``clean`` (``--gittins_clean``, ours) drops
    three more accidents of the reference loop, which is what the real-time controller executes: a waiting
    job is ranked by its attained GPU-time like a running one (the reference ranks it by plain seconds), the
    extra wake-up comes every ``gittins_delta`` (the reference's ``next_unit += now`` makes the period grow
    with the clock), and equal ranks keep running jobs ahead of waiting ones, then submit order."""
    jobs, cluster, log = sim.jobs, sim.cluster, sim.log
    end_events: List[dict] = []
    next_unit = int(jobs.gittins_delta)
    while len(jobs.job_events) + len(jobs.runnable_jobs) > 0:
        if len(jobs.job_events) + len(end_events) == 0:
            sim.print_fn(NOT_LARGE_ENOUGH)
            break
        now, event = _next_event(jobs, end_events, pop_start=False)
        if now > next_unit:
            now, event = next_unit, {}
        for j in event.get("end_jobs", ()):
            cluster.release_job_res(j)
            log.job_complete(j, now)
            jobs.runnable_jobs.remove(j)
        if "start_jobs" in event:
            for j in event["start_jobs"]:
                jobs.move_to_runnable(j)
                j.q_id = 0
            jobs.job_events.pop(0)
        for j in jobs.runnable_jobs:
            _las_clock_update(j, now, gputime, 1, [], jobs.queues, 0)
        if jobs.runnable_jobs:
            att = torch.tensor([_attained(j, gputime) if (j.status == "RUNNING" or clean) else j.executed_time
                                for j in jobs.runnable_jobs], dtype=torch.int64)
            for j, r in zip(jobs.runnable_jobs, gi.gittins_ranks(jobs.job_dist_data, att).tolist()):
                j.rank = r
        if clean:
            sign = -1.0 if descending else 1.0
            jobs.runnable_jobs.sort(key=lambda j: (sign * j.rank, 0 if j.status == "RUNNING" else 1,
                                                   j.submit_time))
        else:
            jobs.runnable_jobs.sort(key=lambda j: j.rank, reverse=descending)
        cluster.empty_infra()
        run_jobs, preempt_jobs = [], []
        for j in jobs.runnable_jobs:
            if cluster.free_gpu >= j.num_gpu:
                if j.status == "PENDING":
                    run_jobs.append(j)
                cluster.free_gpu = int(cluster.free_gpu - j.num_gpu)
            elif j.status == "RUNNING":
                preempt_jobs.append(j)
        _apply_switches(preempt_jobs, run_jobs, now)
        del end_events[:]
        nxt = _earliest_completion(jobs.runnable_jobs, now)
        if nxt is not None:
            end_events.append(nxt)
        if clean:
            next_unit = (now // int(jobs.gittins_delta) + 1) * int(jobs.gittins_delta)
        else:
            next_unit += now
        log.checkpoint(jobs, now)
        sim.mark(now)


# ---- legacy Gandiva ---------------------------------------------------------------------------------------------

def gandiva_legacy(sim: Sim, step: int = 10) -> None:
    """10-second stepping loop over node sets (run_sim.py:101-158, infra/cluster.py:150-489)."""
    jobs, cluster, log = sim.jobs, sim.cluster, sim.log
    if not jobs.job_events:
        return
    cluster.init_gandiva_nodes()
    now = jobs.job_events[0]["time"]
    time_diff = 0
    while len(jobs.job_events) + len(jobs.pending_jobs) + len(jobs.running_jobs) > 0:
        cluster.gandiva_node_set_adjust()
        node_release = cluster.time_slicing_execute(now, jobs, log, time_diff)
        event = next((e for e in jobs.job_events if e["time"] == now), None)
        if event is not None:
            for j in event["start_jobs"]:
                j.executed_time = 0
                if cluster.gandiva_placement(j):
                    j.status = "RUNNING"
                    j.start_time = now
                    jobs.running_jobs.append(j)
                else:
                    jobs.move_to_pending(j)
            jobs.job_events.remove(event)
        if node_release:
            for j in list(jobs.pending_jobs):
                if cluster.gandiva_placement(j):
                    jobs.remove_from_pending(j, now)
                    j.start_time = now
                    jobs.running_jobs.append(j)
        if not jobs.job_events:
            now, time_diff = now + step, step
        else:
            jobs.job_events.sort(key=lambda e: e["time"])
            nxt = jobs.job_events[0]["time"]
            if now + step > nxt:
                time_diff, now = int(nxt - now), nxt
            else:
                now, time_diff = now + step, step
        sim.mark(now)
        if not jobs.job_events and not jobs.running_jobs and jobs.pending_jobs:
            sim.print_fn(NOT_LARGE_ENOUGH)          # nothing runs and nothing will free up
            break


def gpu_demands(sim: Sim) -> None:
    """gpu-demands (run_sim.py:1601-1647): every job runs for its duration from submission; the log
    records the demanded GPUs per job size at every event."""
    jobs, log = sim.jobs, sim.log
    while len(jobs.job_events) + len(jobs.pending_jobs) > 0:
        if not jobs.job_events:
            break
        event = jobs.job_events[0]
        now = event["time"]
        for j in event["end_jobs"]:
            jobs.delete_gpu_job(j)
            j.status = "END"
        for j in event["start_jobs"]:
            j.start_time = j.submit_time
            j.end_time = j.submit_time + j.duration
            j.status = "RUNNING"
            jobs.add_job_end_event(j)
            jobs.add_gpu_job(j)
        jobs.job_events.remove(event)
        jobs.job_events.sort(key=lambda e: e["time"])
        log.checkpoint_gpu_demands(jobs, now)
        sim.mark(now)



# ---- the execution backend's packing / time-slicing policies, simulated ------------------------------------------------

def fifo_shared(sim: Sim, max_share: int = 1, slowdown: Optional[List[float]] = None, quantum: int = 0) -> None:
    """``fifo-pack`` and ``gandiva-slice``: the simulated twins of the real-time controller's ``pack`` mode and
    ``gandiva`` policy (tiresias_b200/exec/controller.py), so that executed and simulated JCTs can be compared.

    The reference models both ideas without executing either: up to 3 tasks share a device and a sharing job runs
    ``1 + FACTOR`` slower (infra/device.py:49-59, infra/node.py:189-205, infra/interference.py:1); Gandiva preempts a
    job every 100 ticks while others wait (core/scheduling/algorithm.py:418-438). Here a job's progress is a RATE:
    a single-GPU job that shares its GPU with k-1 others advances at ``1 / slowdown[k-1]`` (measured table,
    calib/interference.py; ``slowdown = [1, 1.2, 1.2]`` is the reference's constant), everything else at 1. GPUs are
    aligned power-of-two blocks (exec/buddy.py) as in the worker pool; single-GPU jobs prefer a free GPU and only join
    the shared GPU with the fewest tenants when none is left. Waiting jobs are served in list order until the first
    that does not fit (arrivals join the back; the list is kept sorted by GPU count as in ``sim_job_events``).
    With ``quantum`` > 0 (and no packing order change) a job that has held its GPUs for a quantum while others wait
    is preempted and re-queued at the back."""
    from ...exec.buddy import BuddyAllocator, normalize_gpus
    jobs, log = sim.jobs, sim.log
    total = int(sim.cluster.num_gpu)
    pow2 = 1
    while pow2 * 2 <= total:
        pow2 *= 2
    alloc = BuddyAllocator(pow2)
    slow = list(slowdown) if slowdown else [1.0]
    tenants: Dict[int, List[TJob]] = {}
    arrivals = sorted(jobs.job_list, key=lambda j: (j.submit_time, j.job_idx))
    for j in arrivals:
        j.status = "EVENT"
        j.progress = 0.0
    pending: List[TJob] = []
    running: List[TJob] = []
    now = 0.0

    def rate(j: TJob) -> float:
        blk = j.placements[0]["buddy"]
        k = len(tenants.get(blk[0], ())) if (blk[1] == 1 and max_share > 1) else 1
        return 1.0 / slow[min(max(k, 1), len(slow)) - 1]

    def start(j: TJob) -> bool:
        need = normalize_gpus(j.num_gpu, pow2)
        shared = max_share > 1 and need == 1
        blk = alloc.alloc(need, str(j.job_id))
        if blk is None and shared:
            room = [(len(t), g) for g, t in tenants.items() if 0 < len(t) < max_share]
            if room:
                blk = (min(room)[1], 1)
        if blk is None:
            return False
        if shared:
            tenants.setdefault(blk[0], []).append(j)
        j.placements[:] = [{"buddy": list(blk)}]
        j.status = "RUNNING"
        j.slice_start = now
        if j.start_time == MAXSIZE:
            j.start_time = int(round(now))
        running.append(j)
        return True

    def release(j: TJob) -> None:
        blk = tuple(j.placements[0]["buddy"])
        if max_share > 1 and blk[1] == 1 and j in tenants.get(blk[0], ()):
            tenants[blk[0]].remove(j)
            if not tenants[blk[0]]:
                alloc.free(blk)
        else:
            alloc.free(blk)
        running.remove(j)

    while arrivals or pending or running:
        t_arr = arrivals[0].submit_time if arrivals else MAXSIZE
        t_end = min((now + (j.duration - j.progress) / rate(j) for j in running), default=MAXSIZE)
        t_q = MAXSIZE
        if quantum > 0 and pending:
            t_q = min((j.slice_start + quantum for j in running), default=MAXSIZE)
        nxt = min(t_arr, t_end, t_q)
        if nxt >= MAXSIZE:
            sim.print_fn(NOT_LARGE_ENOUGH)
            break
        nxt = max(nxt, now)
        for j in running:
            j.progress += (nxt - now) * rate(j)
        for j in pending:
            j.pending_time = int(round(j.pending_time + (nxt - now)))
        now = nxt
        for j in [j for j in running if j.progress >= j.duration - 1e-6]:
            release(j)
            j.status = "END"
            log.job_complete(j, int(round(now)))
        if quantum > 0 and pending:
            for j in [j for j in running if now - j.slice_start >= quantum - 1e-9]:
                release(j)
                j.status = "PENDING"
                j.preempt = int(j.preempt + 1)
                pending.append(j)
        fresh = []
        while arrivals and arrivals[0].submit_time <= now + 1e-9:
            j = arrivals.pop(0)
            j.start_time = MAXSIZE
            j.pending_time = 0
            fresh.append(j)
        if quantum > 0:
            pending.extend(fresh)
            for j in list(pending):
                if start(j):
                    pending.remove(j)
                else:
                    break
        else:
            for j in pending:                      # removal while iterating skips the next entry (sim_job_events)
                if start(j):
                    pending.remove(j)
                else:
                    break
            for j in fresh:
                if not start(j):
                    j.status = "PENDING"
                    pending.append(j)
            pending.sort(key=lambda j: j.num_gpu)
        sim.mark(int(round(now)))


def dispatch(sim: Sim) -> None:
    """``--schedule`` -> loop (run_sim.py's README-documented names)."""
    global RESUME_OVERHEAD
    RESUME_OVERHEAD = int(getattr(sim.flags, "resume_overhead", 0) or 0)
    s = sim.flags.schedule
    starve = float(getattr(sim.flags, "solve_starvation", 0) or 0)
    if s in ("fifo", "fjf"):
        sim_job_events(sim)
    elif s == "lpjf":
        longest_pending_first(sim)
    elif s in ("sjf", "shortest", "shortest-gpu", "shortest-expected"):
        priority_replace(sim, s)
    elif s == "dlas":
        dlas(sim, gputime=False, solve_starvation=starve)
    elif s in ("dlas-gpu", "dlas-gpu-gittins", "dlas-gpu-pack"):
        dlas(sim, gputime=True, solve_starvation=starve)
    elif s == "multi-dlas-gpu":
        multi_dlas(sim, gputime=True, solve_starvation=starve)
    elif s == "gittins":
        gittins(sim, gputime=True, descending=bool(getattr(sim.flags, "gittins_desc", False)),
                clean=bool(getattr(sim.flags, "gittins_clean", False)))
    elif s == "gandiva-legacy":
        gandiva_legacy(sim)
    elif s == "gpu-demands":
        gpu_demands(sim)
    elif s in ("fifo-pack", "gandiva-slice"):
        table = [float(v) for v in str(getattr(sim.flags, "pack_slowdown", "") or "1,1.2,1.2").split(",")]
        fifo_shared(sim, max_share=int(getattr(sim.flags, "pack_max_share", 3) or 3) if s == "fifo-pack" else 1,
                    slowdown=table, quantum=int(getattr(sim.flags, "slice_quantum", 100) or 100) if s == "gandiva-slice" else 0)
    else:
        raise NotImplementedError(f"unknown schedule {s!r}")


EVENT_SCHEDULES = ("fifo", "fjf", "lpjf", "sjf", "shortest", "shortest-gpu", "shortest-expected", "dlas",
                   "dlas-gpu", "dlas-gpu-gittins", "dlas-gpu-pack", "multi-dlas-gpu", "gittins",
                   "gandiva-legacy", "gpu-demands", "fifo-pack", "gandiva-slice")
