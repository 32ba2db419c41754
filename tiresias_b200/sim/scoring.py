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

UTILCOST = 1.6
MEMCOST = 1.3


def _score(node, task, overflow_term: bool) -> Tuple[Dict[int, float], float, float]:
    device_scores: Dict[int, float] = {}
    total = 0
    best = 999
    for d in node.device_cache.values():
        if not d.can_fit(task):
            continue
        mem_cost = d.get_current_memory() + task.gpu_memory_max / d.memory
        util_cost = d.get_current_utilization()
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
