"""Scheduling policies of the tick simulator (reference: core/scheduling/algorithm.py:182-296,
:418-442; C5).

Signature: ``policy(scheme, placement, infrastructure, jobs_manager, delta, **kw) ->
(nodes, job, success)``; at most one job is placed per call.

``fifo``     looks only at the head of queue 0; if it does not fit, nothing is scheduled this tick.
``horus``    pops up to k=5 jobs (lowest average GPU utilisation first), places the first that fits,
             pushes the rest back.
``horus+``   like horus, but each of the k pops comes from the queue that currently holds the most
             credit, and losers return to the queue they came from.
``gandiva``  fifo plus a post-tick time-slice check: running jobs whose progress hits a multiple of
             the 100-tick quantum are preempted and requeued while anything is waiting (Q13).
"""
from __future__ import annotations

LOOK_AHEAD = 5            # core/scheduling/algorithm.py:204,240
GANDIVA_QUANTUM = 100     # core/scheduling/algorithm.py:426


def schedule_fifo(scheme, placement_algo, infrastructure, jobs_manager, delta, **kwargs):
    job = jobs_manager.get_next_job(delta)
    if job is None:
        return None, None, None
    assert job.is_waiting()
    nodes, success = placement_algo(infrastructure, job, scheme)
    if success:
        jobs_manager.pop(delta)
    return nodes, job, success


def _first_placeable(candidates, placement_algo, infrastructure, scheme):
    for idx, job in enumerate(candidates):
        nodes, success = placement_algo(infrastructure, job, scheme)
        if success:
            return idx, nodes
    return None, None


def schedule_horus(scheme, placement_algo, infrastructure, jobs_manager, delta, k=LOOK_AHEAD, **kwargs):
    count = max(0, min(k, jobs_manager.queuing_jobs(delta)))
    window = []
    for _ in range(count):
        job = jobs_manager.pop(delta)
        assert job.is_waiting()
        window.append(job)
    idx, nodes = _first_placeable(window, placement_algo, infrastructure, scheme)
    if idx is None or nodes is None:
        jobs_manager.insert(window)
        return None, None, False
    winner = window.pop(idx)
    jobs_manager.insert(window)
    return nodes, winner, True


def schedule_horus_plus(scheme, placement_algo, infrastructure, jobs_manager, delta, k=LOOK_AHEAD,
                        **kwargs):
    jq = jobs_manager.job_queue_manager
    before = jobs_manager.total_jobs(delta)
    count = max(0, min(k, jobs_manager.queuing_jobs(delta)))
    if count == 0:
        return None, None, False
    window, origin = [], []
    for _ in range(count):
        jq.update_credits()
        q_idx = jq.queue_idx_by_credit()
        job = jobs_manager.pop(delta, queue_idx=q_idx)
        assert job.is_waiting()
        window.append(job)
        origin.append(q_idx)
    idx, nodes = _first_placeable(window, placement_algo, infrastructure, scheme)
    if idx is None or nodes is None:
        jobs_manager.insert(window, queue_insert_position=origin)
        assert before == jobs_manager.total_jobs(delta)
        return None, None, False
    winner = window.pop(idx)
    origin.pop(idx)
    jobs_manager.insert(window, queue_insert_position=origin)
    assert before - 1 == jobs_manager.total_jobs(delta)
    return nodes, winner, True


def time_slice_check(infrastructure, jobs_manager, quanta: int = GANDIVA_QUANTUM) -> None:
    """Gandiva's suspend-resume: only when something is queued."""
    if jobs_manager.queuing_jobs() == 0:
        return
    due = []
    for job_id, job in jobs_manager.running_jobs.items():
        done = job.time_processed()
        if done > 1 and done % quanta == 0:
            due.append(job_id)
    for job_id in due:
        jobs_manager.preempt(job_id, infrastructure)


scheduling_algorithms = {
    "fifo": schedule_fifo,
    "horus": schedule_horus,
    "horus+": schedule_horus_plus,
    "gandiva": schedule_fifo,
}

plugin_algorithms = {"gandiva": time_slice_check}
