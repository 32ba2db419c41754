"""Placement algorithms of the tick simulator (reference: core/scheduling/algorithm.py:15-180,
:299-415; C6).

``yarn``   first free node with enough idle devices / CPU / memory; jobs larger than a node walk the
           nodes in id order filling workers, all-or-nothing with rollback (needs at least
           ``ceil(gpus / gpus_per_node)`` nodes).
``horus``  (also ``horus+``, ``gandiva``) score every free node per task, keep the ``gpu_demand`` best
           in a bounded heap, grow a trial plan from each candidate outward by rack distance, undo
           it, then commit the plan that spans the fewest nodes. Devices may be shared (<= 3 tasks,
           500 MiB margin).

Every function takes ``(infrastructure, job, scheme)`` and returns ``(nodes_by_id, success)``; on
success the tasks are reserved (``placed_tasks``) and ``job.tasks_running_on`` is filled in.
Order-sensitive details that decide ties — heap layout, stable sorts, dict insertion order, and the
reference's partial clean-up of trial plans (``placed_jobs`` survives on all but the first node of a
trial) — are kept because ``cluster.csv`` parity depends on them.
"""
from __future__ import annotations

import math
from heapq import heappop, heappush
from typing import Dict, List, Tuple

from . import util
from .scoring import score_fn


class NodeDeviceInfo:
    """Heap entry: ``__lt__`` is inverted so the heap's root is the *worst* (largest) best-device
    score and trimming the heap drops the worst candidate."""

    __slots__ = ("node", "device_scores", "total_score", "min_score")

    def __init__(self, node, device_scores, total_score, min_score):
        self.node = node
        self.device_scores = device_scores
        self.total_score = total_score
        self.min_score = min_score

    def __lt__(self, other) -> bool:
        return self.min_score > other.min_score


# ---- yarn ------------------------------------------------------------------------------------------

def ms_yarn_placement(infrastructure, next_job, scheme=None):
    if next_job.gpus > infrastructure.num_gpu_p_node:
        return try_cross_node_alloc_ms(infrastructure, next_job)
    return try_single_node_alloc_ms(infrastructure, next_job)


def try_single_node_alloc_ms(infrastructure, job):
    need_cpu, need_mem = job.total_cpus_required(), job.total_mem_required()
    for node in infrastructure.get_free_nodes():
        if (node.get_free_devices(False) >= job.gpus and node.cpu_free() >= need_cpu
                and node.mem_free() >= need_mem and node.try_alloc_job(job, True)):
            return {node.node_id: node}, True
    return {}, False


def try_cross_node_alloc_ms(infrastructure, job, sort_fn=None, filter_fn=None):
    """Workers are handed to nodes in id order. Per node up to ``can_fit + 1`` workers are tried
    (the reference's ``<=``); a worker the node then refuses ends that node's turn."""
    least_nodes = math.ceil(job.gpus / infrastructure.num_gpu_p_node)
    total = len(job.tasks)
    waiting = dict(job.tasks)                  # tasks not yet assigned (all alike)
    assigned: Dict[str, object] = {}
    nodes_assigned: Dict[str, object] = {}
    nodes = infrastructure.nodes.values()
    if filter_fn:
        nodes = filter_fn(nodes)
    if sort_fn:
        nodes = sort_fn(nodes)

    for node in nodes:
        if not node.is_free():
            continue
        if len(assigned) == total:
            break
        room = node.can_fit_num_task(waiting)
        if room == 0:
            continue
        taken = 0
        refused = False
        for task_id, task in job.tasks.items():
            if task_id in assigned or "worker" not in task_id or taken > room:
                continue
            waiting.pop(task_id, None)
            taken += 1
            if not node.try_reserve_and_placed_task(task):
                waiting[task_id] = task
                taken -= 1
                refused = True
                break
            job.tasks_running_on[task_id] = node.node_id
            assigned[task_id] = task
        if taken > 0:
            node.try_reserve_and_placed_job(job, False)
            nodes_assigned[node.node_id] = node
        if refused:
            continue
        if len(nodes_assigned) >= least_nodes and total == len(assigned):
            break

    if len(assigned) < total or len(nodes_assigned) < least_nodes:
        for node in nodes_assigned.values():
            node.placed_jobs.pop(job.job_id)
            for t in job.tasks.values():
                held = node.placed_tasks.pop(t.task_id, None)
                if held is not None:
                    node.release_allocated_resources(held, reserved=True)
        return {}, False
    if infrastructure.verbose:
        util.print_fn("placed job %s with task %d, on node - %s"
                      % (job.job_id, job.worker_count, str(nodes_assigned.keys())))
    return nodes_assigned, True


# ---- horus family ------------------------------------------------------------------------------------

def _candidates(infrastructure, job, scheme) -> List[NodeDeviceInfo]:
    """Best ``gpu_demand`` (node, task) scorings, best first (stable on the heap's array order).

    A node whose devices hold no task scores every device alike and consumes no utilisation draw
    (sim/scoring.py fast path), so the verdict of ``can_fit`` and the scores depend only on the node's free CPU /
    memory, its device memories and the task: those are memoised per call (most of a 4x32x8 cluster is such nodes;
    the reference recomputes them per node, 90 000 scorings on a 300-job trace)."""
    score = score_fn[scheme]
    demand = job.gpus
    heap: List[NodeDeviceInfo] = []
    memo: Dict[tuple, tuple] = {}
    for t in job.tasks.values():
        tkey = (t.cpu, t.mem, t.gpu, t.gpu_memory_max, t.gpu_utilization_max)
        for node in infrastructure.get_free_nodes():
            if node.devices_empty() and not node.verbose:
                key = (tkey, node.cpu_free(), node.mem_free(), node.gpu_memory_capacity, len(node.device_cache))
                hit = memo.get(key)
                if hit is None:
                    hit = (score(node, t)[1:] if node.can_fit(t, pack=True) else None,)
                    memo[key] = hit
                if hit[0] is None:
                    continue
                node_sum, best = hit[0]
                heappush(heap, NodeDeviceInfo(node, None, node_sum, best))
            else:
                if not node.can_fit(t, pack=True):
                    continue
                device_scores, node_sum, best = score(node, t)
                heappush(heap, NodeDeviceInfo(node, device_scores, node_sum, best))
            if len(heap) > demand:
                heappop(heap)
    return sorted(heap, key=lambda info: info.min_score)


def _grow_plan(infrastructure, job, start_node) -> Tuple[set, Dict[str, str]]:
    """Trial plan seeded at ``start_node``: everything that fits there, then the remaining tasks
    over the racks by distance (nodes in rack order, the seed node included again)."""
    nodes_used: set = set()
    where: Dict[str, str] = {}
    total = len(job.tasks)

    def offer(node, task_id, task) -> None:
        if node.try_reserve_and_placed_task(task, pack=True):
            assert node.try_reserve_and_placed_job(job, False)
            nodes_used.add(node.node_id)
            if task_id in where:
                raise ValueError(f"task {task_id} placed twice")
            where[task_id] = node.node_id

    for task_id, task in job.tasks.items():
        offer(start_node, task_id, task)
    for rack_id, _ in infrastructure.get_racks_by_dist(start_node.rack_id):
        if len(where) >= total:
            break
        for node in infrastructure.racks[rack_id].nodes.values():
            if len(where) >= total:
                break
            for task_id, task in job.tasks.items():
                if task_id in where:
                    continue
                offer(node, task_id, task)
                if len(where) >= total:
                    break
    return nodes_used, where


def _undo_plan(infrastructure, job, where: Dict[str, str]) -> None:
    """Release a trial plan. Only the node of the plan's *first* task forgets the job; the other
    nodes keep it in ``placed_jobs`` (reference behaviour, core/scheduling/algorithm.py:125-136)."""
    first = True
    for task_id, node_id in where.items():
        node = infrastructure.nodes[node_id]
        if first:
            node.placed_jobs.pop(job.job_id)
            first = False
        held = node.placed_tasks.pop(task_id, None)
        if held is not None:
            node.release_allocated_resources(held, reserved=True)


def horus_placement(infrastructure, next_job, scheme):
    total = len(next_job.tasks)
    plans: List[Tuple[set, Dict[str, str]]] = []
    for info in _candidates(infrastructure, next_job, scheme):
        node = infrastructure.nodes[info.node.node_id]
        nodes_used, where = _grow_plan(infrastructure, next_job, node)
        _undo_plan(infrastructure, next_job, where)
        if len(where) >= total:
            plans.append((nodes_used, where))
    if not plans:
        return None, False
    plans.sort(key=lambda p: len(p[0]))                 # fewest nodes; stable
    where = plans[0][1]
    results = {}
    placed = 0
    for task_id, node_id in where.items():
        node = infrastructure.nodes[node_id]
        if node.try_reserve_and_placed_task(next_job.tasks[task_id], pack=True):
            results[node_id] = node
            next_job.tasks_running_on[task_id] = node.node_id
            node.try_reserve_and_placed_job(next_job, False)
            placed += 1
    assert placed == total, "has %d expected %d - assigned map %s " % (placed, total, where)
    return results, True


placement_algorithms = {
    "yarn": ms_yarn_placement,
    "horus": horus_placement,
    "horus+": horus_placement,
    "gandiva": horus_placement,
}
