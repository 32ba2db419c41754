"""Device scoring for horus / horus+ / gandiva placement (reference: core/scheduling/horus.py:1-49,
C7).

Per device that can still take the task:
``cost = (cur_mem + task_mem/dev_mem) * 1.3 + util_term/100 + n_resident`` where, for horus,
``util_term = |cur_util + task.util_max - 100|`` (x1.6 when above 100) and, for gandiva, just
``cur_util``. ``cur_mem + task_mem/dev_mem`` is the reference's operator precedence (absolute MiB
plus a fraction), kept on purpose. ``cur_util`` consumes one Gaussian draw per resident task.
"""
from __future__ import annotations

from typing import Dict, Tuple

from .infra import MEMORY_MARGIN_MIB, sample_utilizations

UTILCOST = 1.6
MEMCOST = 1.3


def _score(node, task, overflow_term: bool) -> Tuple[Dict[int, float], float, float]:
    device_scores: Dict[int, float] = {}
    total = 0
    best = 999
    devices = node.device_cache.values()
    if not any(d.running_tasks for d in devices):
        # Fast path for a node with nothing resident: every device scores the same and no utilisation
        # draw is consumed (the reference draws once per *resident* task), so one evaluation serves
        # all devices; the sum is accumulated device by device to stay bit-identical.
        for d in devices:
            if not d.memory - (0 + task.gpu_memory_max) > MEMORY_MARGIN_MIB:
                return device_scores, total, best
            mem_cost = 0 + task.gpu_memory_max / d.memory
            util_cost = 0.0
            if overflow_term:
                util_cost = (util_cost + task.gpu_utilization_max) - 100.0
                util_cost = abs(util_cost) * UTILCOST if util_cost > 0 else abs(util_cost)
            cost = (mem_cost * MEMCOST) + (util_cost / 100) + 0
            break
        for d in devices:
            device_scores[d.device_id] = cost
            total += cost
        return device_scores, total, cost
    fitting = [d for d in devices if d.can_fit(task)]
    utils = sample_utilizations(fitting, fitting[0].rng) if fitting else []
    for d, (util_cost, _) in zip(fitting, utils):
        mem_cost = d.get_current_memory() + task.gpu_memory_max / d.memory
        if overflow_term:
            util_cost = (util_cost + task.gpu_utilization_max) - 100.0
            util_cost = abs(util_cost) * UTILCOST if util_cost > 0 else abs(util_cost)
        cost = (mem_cost * MEMCOST) + (util_cost / 100) + len(d.running_tasks)
        device_scores[d.device_id] = cost
        total += cost
        if cost < best:
            best = cost
    return device_scores, total, best


def horus_score(node, task):
    return _score(node, task, True)


def gandiva_score(node, task):
    return _score(node, task, False)


score_fn = {"horus": horus_score, "horus+": horus_score, "gandiva": gandiva_score}
