"""Network-balanced placement (reference: core/lp.py:2-242, D14).

The reference states a MathProg model — binary ``p[j,n]`` places component ``j`` (PS shard or
worker) on node ``n``; minimise the maximum per-node network load subject to free GPUs / CPUs and
every component placed exactly once — but its solver import (CPLEX) is commented out and
``placement()`` is ``pass``. Here the same model is solved with SciPy's MILP front-end (HiGHS) when
the instance is small enough, and by a longest-processing-time greedy otherwise; both honour the
same constraints. ``prepare_job_info`` / ``prepare_cluster_info`` keep the reference's input layout.
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np

from .cluster import PS_CPU, WORKER_CPU
from .jobs import TJob

MILP_VAR_LIMIT = 48           # exact solve only for small instances (<= 4 workers on a few nodes)


def prepare_job_info(job: TJob) -> Dict:
    num_ps, num_w = len(job.ps_network), job.num_gpu
    assert num_w == 1 or num_ps == num_w
    return {"t": [round(ps, 1) for ps in job.ps_network], "num_ps": num_ps, "num_w": num_w,
            "m_size": job.w_network[0], "ps_c": PS_CPU, "w_c": WORKER_CPU}


def prepare_cluster_info(cluster) -> Dict:
    nodes = list(cluster.nodes())
    return {"fg": [n.free_gpus for n in nodes], "fc": [n.free_cpus for n in nodes],
            "ct": [n.network_out for n in nodes], "num_n": len(nodes)}


def _solve_milp(job_info: Dict, cl: Dict) -> Optional[np.ndarray]:
    """x[j, n] for components j = ps_0..ps_{P-1}, w_0..w_{W-1}; returns the 0/1 matrix or None."""
    try:
        from scipy.optimize import Bounds, LinearConstraint, milp
    except Exception:
        return None
    P, W, N = job_info["num_ps"], job_info["num_w"], cl["num_n"]
    J = P + W
    if J * N + 1 > MILP_VAR_LIMIT:
        return None
    g = np.array([0] * P + [1] * W, dtype=float)
    c = np.array([PS_CPU] * P + [WORKER_CPU] * W, dtype=float)
    t = np.array([job_info["t"][i] * W for i in range(P)] + [job_info["m_size"]] * W, dtype=float)
    nv = J * N + 1                                   # last variable = maxt
    cost = np.zeros(nv)
    cost[-1] = 1.0
    rows, lo, hi = [], [], []
    for n in range(N):                               # ct[n] + sum_j p[j,n] t[j] <= maxt
        r = np.zeros(nv)
        r[n:J * N:N] = t
        r[-1] = -1.0
        rows.append(r); lo.append(-np.inf); hi.append(-cl["ct"][n])
        r = np.zeros(nv); r[n:J * N:N] = g
        rows.append(r); lo.append(-np.inf); hi.append(cl["fg"][n])
        r = np.zeros(nv); r[n:J * N:N] = c
        rows.append(r); lo.append(-np.inf); hi.append(cl["fc"][n])
    for j in range(J):                               # every component exactly once
        r = np.zeros(nv); r[j * N:(j + 1) * N] = 1.0
        rows.append(r); lo.append(1.0); hi.append(1.0)
    integrality = np.ones(nv); integrality[-1] = 0
    ub = np.ones(nv); ub[-1] = np.inf
    res = milp(cost, constraints=LinearConstraint(np.array(rows), lo, hi), integrality=integrality,
               bounds=Bounds(np.zeros(nv), ub), options={"time_limit": 0.25})
    if not res.success or res.x is None:
        return None
    return np.rint(res.x[:-1]).reshape(J, N)


def _solve_greedy(job_info: Dict, cl: Dict, consolidate: bool) -> Optional[np.ndarray]:
    P, W, N = job_info["num_ps"], job_info["num_w"], cl["num_n"]
    fg, fc, load = list(cl["fg"]), list(cl["fc"]), list(cl["ct"])
    x = np.zeros((P + W, N))
    comps = [(job_info["t"][i] * W, i, 0, PS_CPU) for i in range(P)] + \
            [(job_info["m_size"], P + i, 1, WORKER_CPU) for i in range(W)]
    comps.sort(key=lambda c: -c[0])
    used_nodes = set()
    for traffic, j, need_g, need_c in comps:
        cand = [n for n in range(N) if fg[n] >= need_g and fc[n] >= need_c]
        if not cand:
            return None
        key = (lambda n: (n not in used_nodes, load[n])) if consolidate else (lambda n: load[n])
        n = min(cand, key=key)
        x[j, n] = 1
        fg[n] -= need_g; fc[n] -= need_c; load[n] += traffic
        used_nodes.add(n)
    return x


def parse_lp_solution(cluster, job: TJob, x: np.ndarray, job_info: Dict) -> None:
    """Turn the 0/1 matrix into node allocations + ``job.placements`` (one record per node)."""
    P, W = job_info["num_ps"], job_info["num_w"]
    nodes = list(cluster.nodes())
    for n_idx, node in enumerate(nodes):
        local_w = int(x[P:, n_idx].sum())
        local_ps = [i for i in range(P) if x[i, n_idx] > 0.5]
        num_cpu = local_w * WORKER_CPU + len(local_ps) * PS_CPU
        if num_cpu == 0:
            continue
        load = round(local_w * job_info["m_size"], 1)
        for i in local_ps:
            load += round(job_info["t"][i] * (W - 2 * local_w), 1)
        load = round(max(load, 0.0), 1)
        node.alloc_job_res(local_w, num_cpu)
        node.add_network_load(load, load)
        job.placements.append({"switch": n_idx // cluster.num_node_p_switch, "nodes": [
            {"id": n_idx % cluster.num_node_p_switch, "num_gpu": local_w, "num_cpu": num_cpu, "tasks": [],
             "network": load}]})


def placement(cluster, job: TJob, consolidate: bool = False) -> bool:
    info, cl = prepare_job_info(job), prepare_cluster_info(cluster)
    if info["num_w"] == 1:
        x = _solve_greedy(info, cl, True)
    else:
        x = None if consolidate else _solve_milp(info, cl)
        if x is None:
            x = _solve_greedy(info, cl, consolidate)
    if x is None:
        return False
    parse_lp_solution(cluster, job, x, info)
    return True
