"""Network cost of a placed distributed job (reference: core/network/network_service.py:3-39, C12;
per-node traffic model infra/switch.py:97-133, D12).

``analytic``   the reference's formula, made executable:
               ``extra = (model_size/bandwidth + cross_nodes*latency) * iterations * 2``
               (2 = push + pull round trip). In the reference this path raises because its live
               ``Job`` has no ``ps_count`` / ``model_size`` / ``iterations`` and never creates a PS
               task; here the placement is read the way infra/switch.py lays jobs out —
               every GPU hosts one worker *and* one PS shard — so ``cross_nodes`` is the number of
               nodes the job spans when it spans more than one.
``calibrated`` per-iteration push+pull seconds measured on the 8xB200 box by
               :mod:`tiresias_b200.calib` (fused in-kernel reduce-scatter / all-gather over NVSwitch),
               keyed by model and worker count, times ``iterations``.
"""
from __future__ import annotations

from typing import Dict, Optional

from . import util
from .models import get_model_size

ROUND_TRIP = 2.0


def crossed_nodes(job) -> int:
    """Number of nodes whose PS shards and workers talk across the network."""
    ps_nodes, wk_nodes = set(), set()
    for task_id, node_id in job.tasks_running_on.items():
        (ps_nodes if "ps" in task_id else wk_nodes).add(node_id)
    if not ps_nodes:                       # worker-only jobs: one PS shard per worker GPU
        return len(wk_nodes) if len(wk_nodes) > 1 else 0
    return len(ps_nodes.symmetric_difference(wk_nodes))


def calculate_network_costs(infrastructure, job, calibration: Optional[Dict] = None) -> float:
    """Seconds to add to ``job.duration`` for its gradient/parameter traffic."""
    if not job.is_distributed():
        return 0
    iterations = job.iterations or 0
    if calibration is not None:
        per_iter = lookup_comm_seconds(calibration, job.model_name, int(job.gpus))
        if per_iter is not None:
            return per_iter * iterations
    cross = crossed_nodes(job)
    if cross == 0:
        return 0
    size = job.model_size if job.model_size is not None else get_model_size(job.model_name or "")
    model_per_sec = size / infrastructure.bandwidth
    nodes_induced_sec = cross * infrastructure.internode_latency
    extra = (model_per_sec + nodes_induced_sec) * (iterations * ROUND_TRIP)
    util.print_fn("Cross %d nodes need to added extra %f for job %s" % (cross, extra, job.job_id))
    return extra


def lookup_comm_seconds(calibration: Dict, model_name: Optional[str], gpus: int) -> Optional[float]:
    """``calibration["comm_s_per_iter"][model][str(gpus)]`` with nearest-lower GPU count fallback."""
    table = (calibration or {}).get("comm_s_per_iter", {}).get(model_name or "")
    if not table:
        return None
    keys = sorted(int(k) for k in table)
    pick = max([k for k in keys if k <= gpus], default=keys[0])
    return float(table[str(pick)])


def node_traffic_mb(model_size_mb: float, ps_shard_mb, local_gpus: int, total_gpus: int) -> float:
    """Per-iteration MB through one node's NIC under the one-worker-one-PS-shard-per-GPU layout
    (infra/switch.py:97-107): local workers move the whole model, each local PS shard exchanges its
    shard with every non-local worker, minus what stays inside the node."""
    traffic = model_size_mb * local_gpus
    for shard in ps_shard_mb:
        traffic += shard * (total_gpus - local_gpus)
        traffic -= shard * local_gpus
    return round(traffic, 1)
