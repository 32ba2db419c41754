"""Cluster state: switch/rack -> node -> GPU device (reference: infra/infrastructure.py:16-146,
infra/rack.py:5-16, infra/node.py:7-287, infra/device.py:4-58, infra/interference.py:1; C11).

Parity notes: node ids are global 1-based strings, rack ids 0-based strings; a node is
"free" when it has CPU *or* memory left (Q9); a device takes at most 3 co-located tasks and keeps a
500 MiB margin (Q22); a node counts as idle only until it first hosts a job because ``placed_jobs``
is never cleared on completion (Q21); utilisation is one Gaussian draw per resident task from the
legacy global NumPy generator (Q14). All of that is kept bit for bit; the calibrated mode swaps the
interference constant and the utilisation draw for measured B200 tables (``calibration``).
"""
from __future__ import annotations

import csv
import os
from collections import OrderedDict
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np

from . import util

FACTOR = 0.2                      # infra/interference.py:1
MAX_TASKS_PER_DEVICE = 3          # infra/device.py:53-55 ("PCIe bandwidth setting")
MEMORY_MARGIN_MIB = 500           # infra/device.py:57
RESOURCE_INSUFFICIENT_MSG = "Node %s - %s not enough for task %s, needed %d, got %d free"  # infra/msg.py

SPEC_KEYS = ("num_switch", "num_node_p_switch", "num_gpu_p_node", "num_cpu_p_node", "mem_p_node")


def global_rng() -> np.random.RandomState:
    """The legacy global generator behind ``np.random.normal`` / ``np.random.seed``."""
    return np.random.mtrand._rand


class MeanRandomState:
    """``--fast``: stands in for the legacy generator wherever utilisation is drawn — every Gaussian is
    its mean — and hands every other request (k-means' ``randint`` ...) to the real generator, so nothing
    else changes. SURVEY 7.4 item 2: the exact mode replays the reference's stream, this one trades
    column 6 of cluster.csv for not touching it."""

    def __init__(self, real: np.random.RandomState):
        self._real = real

    def standard_normal(self, size=None):
        return 0.0 if size is None else np.zeros(size)

    def normal(self, loc=0.0, scale=1.0, size=None):
        return loc if size is None and np.ndim(loc) == 0 else np.broadcast_to(np.asarray(loc, dtype=np.float64),
                                                                             np.shape(loc) if size is None else size).copy()

    def __getattr__(self, name):
        return getattr(self._real, name)


class Device:
    """One GPU. ``running_tasks`` holds every task *placed* on it (reserved or running)."""

    __slots__ = ("device_id", "node_id", "memory", "enable_pack", "running_tasks", "rng", "_normal", "_mem",
                 "_ndraw")

    def __init__(self, device_id, node_id, memory_cap, enable_pack, rng=None):
        self.device_id = device_id
        self.node_id = node_id
        self.memory = memory_cap
        self.enable_pack = enable_pack
        self.running_tasks: "OrderedDict[str, object]" = OrderedDict()
        self.rng = rng if rng is not None else global_rng()
        self._normal = self.rng.normal
        self._mem = 0                   # cached get_current_memory(); None = recompute
        self._ndraw = 0                 # resident tasks whose utilisation is drawn (not measured)

    def is_idle(self) -> bool:
        return not self.running_tasks

    def sample_utilization(self) -> Tuple[float, bool]:
        """(value, came_from_a_draw). One N(avg, (max-avg)/2) draw per resident task, each capped
        at 100, running sum capped at 100. The flag records whether the reference's value would be
        a 1-element ndarray (any uncapped draw contributed) or a plain int — that decides how
        ``cluster.csv`` prints it (``[4.51]`` vs ``100``)."""
        if not self.running_tasks:
            return 0.0, False
        total, drawn = 0.0, False
        normal = self._normal
        for t in self.running_tasks.values():
            if t.measured_profile:
                total += min(t.gpu_utilization_avg, 100)   # calibrated mode: the measured value, no draw (K7)
                if 100 < total:
                    total, drawn = 100.0, False
                continue
            x = normal(t.gpu_utilization_avg, (t.gpu_utilization_max - t.gpu_utilization_avg) / 2)
            if x < 100:
                total += x
                drawn = True
            else:
                total += 100
            if 100 < total:
                total, drawn = 100.0, False
        return total, drawn

    def get_current_utilization(self) -> float:
        return self.sample_utilization()[0]

    def get_current_memory(self):
        """Sum of the resident tasks' peak memory, each and the total capped at the capacity. Cached
        between changes of the resident set (recomputed in resident order, so the float sum is the
        reference's)."""
        mem = self._mem
        if mem is None:
            mem = 0
            cap = self.memory
            for t in self.running_tasks.values():
                m = t.gpu_memory_max
                mem += cap if cap < m else m
                if cap < mem:
                    mem = cap
            self._mem = mem
        return mem

    def can_fit(self, task) -> bool:
        if len(self.running_tasks) >= MAX_TASKS_PER_DEVICE:
            return False
        return self.memory - (self.get_current_memory() + task.gpu_memory_max) > MEMORY_MARGIN_MIB

    def add_task(self, task, pack: bool = False) -> bool:
        if not self.can_fit(task):
            return False
        if not pack and self.running_tasks:
            return False
        if task.task_id not in self.running_tasks and not task.measured_profile:
            self._ndraw += 1
        self.running_tasks[task.task_id] = task
        self._mem = None
        return True

    def pop_task(self, task_id):
        self._mem = None
        task = self.running_tasks.pop(task_id, None)
        if task is not None and not task.measured_profile:
            self._ndraw -= 1
        return task

    def reset(self) -> None:
        self.running_tasks = OrderedDict()
        self._mem = 0
        self._ndraw = 0


_IDLE = (0.0, False)


def sample_utilizations(devices, rng) -> List[Tuple[float, bool]]:
    """:meth:`Device.sample_utilization` for several devices with ONE call into the generator. The
    legacy ``RandomState.normal(loc, scale)`` is ``loc + scale * gauss()`` on the same Gaussian stream
    that ``standard_normal(n)`` fills element by element, so drawing all resident tasks' Gaussians at once
    (device order, then resident order) and applying loc / scale here consumes the stream and produces
    the doubles exactly as the reference's one-call-per-task does — at a fraction of the overhead (this
    function is the simulator's hottest: horus scores every candidate node with it)."""
    n = 0
    for d in devices:
        n += d._ndraw
    gs = rng.standard_normal(n).tolist() if n else ()
    out = []
    i = 0
    for d in devices:
        tasks = d.running_tasks
        if not tasks:
            out.append(_IDLE)
            continue
        total, drawn = 0.0, False
        for t in tasks.values():
            if t.measured_profile:
                total += min(t.gpu_utilization_avg, 100)   # calibrated mode (K7): measured, no draw
                if 100 < total:
                    total, drawn = 100.0, False
                continue
            avg = t.gpu_utilization_avg
            x = avg + ((t.gpu_utilization_max - avg) / 2) * gs[i]
            i += 1
            if x < 100:
                total += x
                drawn = True
            else:
                total += 100
            if 100 < total:
                total, drawn = 100.0, False
        out.append((total, drawn))
    return out


class Node:
    """A server: CPU / memory counters plus its devices and the tasks/jobs placed on it."""

    def __init__(self, rack_id, node_id, gpu_memory_capacity=0, cpus=0, gpus=0, memory=0,
                 enable_pack=False, rng=None, verbose=False):
        self.node_id = str(node_id)
        self.rack_id = rack_id
        self.cpu_count, self.gpu_count, self.mem_size = cpus, gpus, memory
        self.gpu_memory_capacity = gpu_memory_capacity
        self.network_usage = 0
        self.cpu_used = self.gpu_used = self.mem_used = 0
        self.enable_pack = enable_pack
        self.verbose = verbose
        self.device_cache: "OrderedDict[int, Device]" = OrderedDict(
            (i, Device(i, self.node_id, gpu_memory_capacity, enable_pack, rng)) for i in range(gpus))
        self.running_tasks: Dict[str, object] = {}
        self.placed_tasks: Dict[str, object] = {}
        self.placed_jobs: Dict[str, object] = {}
        self.finished_tasks: List[str] = []
        self.interference_factor = FACTOR
        # tasks resident on this node's devices, summed over devices (every device mutation goes through the three
        # Node methods below): `devices_empty()` lets the placement memoise the scoring of untouched nodes
        self._dev_tasks = 0

    # -- counters ---------------------------------------------------------------------------------
    def get_network_usage(self):
        return self.network_usage

    def check_util(self) -> Tuple[float, float]:
        return float(self.cpu_used) / float(self.cpu_count), float(self.mem_used) / float(self.mem_size)

    def resize_node(self, cpus, gpus, memory) -> None:
        self.cpu_count, self.gpu_count, self.mem_size = cpus, gpus, memory

    def cpu_free(self):
        return self.cpu_count - self.cpu_used

    def mem_free(self):
        return self.mem_size - self.mem_used

    def gpu_free(self) -> int:
        return self.get_free_devices(False)

    def is_free(self) -> bool:
        return self.cpu_count - self.cpu_used > 0 or self.mem_size - self.mem_used > 0

    def is_idle(self) -> bool:
        return not (self.running_tasks or self.placed_tasks or self.placed_jobs)

    def devices_empty(self) -> bool:
        return self._dev_tasks == 0

    def get_free_devices(self, pack: bool) -> int:
        if pack:
            return len(self.device_cache)
        return sum(1 for d in self.device_cache.values() if not d.running_tasks)

    def reset_resource(self, num=0, gpu=1, cpu=4, mem=6) -> None:
        """Legacy helper (infra/node.py:62-69; the reference's version cannot run because it
        iterates the device dict without ``.items()``)."""
        assert len(self.running_tasks) >= num
        self.cpu_used, self.mem_used = num * cpu, num * mem
        for d in self.device_cache.values():
            d.reset()
        self._dev_tasks = 0

    # -- capacity questions -------------------------------------------------------------------------
    def can_fit_num_task(self, tasks, pack: bool = False) -> int:
        """How many of ``tasks`` (all alike) fit: min over the GPU, CPU and memory head-room, each
        clipped to ``len(tasks)`` (may go negative when a resource is over-committed)."""
        if pack:
            return self._can_fit_num_with_pack(tasks)
        first = next(iter(tasks.values()))
        n = len(tasks)

        def clip(capacity) -> int:
            slack = capacity - n
            return n if slack >= 0 else n + slack

        by_gpu = clip(self.get_free_devices(False) // first.gpu)
        by_cpu = clip(self.cpu_free() // first.cpu)
        by_mem = clip(self.mem_free() // first.mem)
        return min(min(by_cpu, by_mem), by_gpu)

    def _can_fit_num_with_pack(self, tasks) -> int:
        count = 0
        for t in tasks.values():
            if self.cpu_free() < t.cpu or self.mem_free() < t.mem:
                continue
            for d in self.device_cache.values():
                if d.memory - (d.get_current_memory() + t.gpu_memory_max) > MEMORY_MARGIN_MIB:
                    count += 1
        return count

    def can_fit(self, task, pack: bool = False) -> bool:
        cpu_left = self.cpu_free() - task.cpu
        mem_left = self.mem_free() - task.mem
        if cpu_left < 0 or mem_left < 0:
            if self.verbose:
                kind, need, got = ("cpu", task.cpu, self.cpu_free()) if cpu_left < 0 else \
                    ("mem", task.mem, self.mem_free())
                util.logger.info(RESOURCE_INSUFFICIENT_MSG, self.node_id, kind, task.task_id, need, got)
            return False
        if not pack:
            return self.get_free_devices(False) - task.gpu >= 0
        fits = False
        for d in self.device_cache.values():       # every device is probed, as in the reference
            if not d.running_tasks:                # (an empty device only has the memory margin to pass)
                if d.memory - (0 + task.gpu_memory_max) > MEMORY_MARGIN_MIB:
                    fits = True
            elif d.can_fit(task):
                fits = True
        return fits

    # -- reservation ----------------------------------------------------------------------------------
    def try_reserve_and_placed_task(self, task, pack: bool = False) -> bool:
        """Reserve CPU/memory and ``task.gpu`` devices (walked in index order). When fewer devices
        accept the task than it needs, the call reports failure but — as in the reference
        (infra/node.py:212-233) — keeps the partial reservation; callers never undo it."""
        if not self.can_fit(task, pack=pack):
            return False
        self.cpu_used += task.cpu
        self.mem_used += task.mem
        missing = task.gpu
        for d in self.device_cache.values():
            if missing <= 0:
                break
            new = task.task_id not in d.running_tasks
            if d.add_task(task, pack=pack):
                missing -= 1
                if new:
                    self._dev_tasks += 1
        if missing == 0:
            self.placed_tasks[task.task_id] = task
        return missing == 0

    def try_reserve_and_placed_job(self, job, count_task_in_current_node: bool = False) -> bool:
        """Register ``job`` on this node. With ``count_task_in_current_node`` the job's *first*
        task must already be placed here, otherwise nothing is registered."""
        found = False
        for t in job.tasks.keys():
            if t in self.placed_tasks:
                found = True
                break
            if count_task_in_current_node:
                return False
        self.placed_jobs[job.job_id] = job
        return found

    def try_alloc_job(self, job, is_single: bool = False) -> bool:
        """Place every task of ``job`` on this node (single-node yarn path)."""
        result = False
        if self.can_fit_num_task(job.tasks) >= job.task_count:
            placed = 0
            for t in list(job.tasks.values()):
                result = self.try_reserve_and_placed_task(t)
                if result:
                    job.tasks_running_on[t.task_id] = self.node_id
                    placed += 1
            if placed > 0:
                result = self.try_reserve_and_placed_job(job, is_single)
                if not result:
                    # the reference's undo of this case cannot run (infra/node.py:276-279 iterates
                    # dict items as tasks); it is only reached when the first task failed to place
                    # while a later one succeeded. Undo properly and report failure.
                    for t in job.tasks.values():
                        job.tasks_running_on.pop(t.task_id, None)
                        if self.placed_tasks.pop(t.task_id, None) is not None:
                            self.release_allocated_resources(t, reserved=True)
                    self.placed_jobs.pop(job.job_id, None)
                    util.print_fn("RELEASED: Job does not fit on node", util.LOG_LEVEL_WARNING)
                    return result
                if self.verbose:
                    util.print_fn("placed SINGLE NODE job %s, num tasks %d on node %s"
                                  % (job.job_id, len(job.tasks), self.node_id))
        elif self.verbose:
            util.print_fn("Job does not fit on node", util.LOG_LEVEL_WARNING)
        return result

    def release_allocated_resources(self, task, reserved: bool = False) -> Dict[str, str]:
        """Give back the task's CPU/memory and take it off every device. Returns the tasks that are
        now alone on their device and were marked interfered ({task_id: job_id})."""
        self.cpu_used -= task.cpu
        self.mem_used -= task.mem
        relieved: Dict[str, str] = {}
        for d in self.device_cache.values():
            if task.task_id in d.running_tasks:
                d.pop_task(task.task_id)
                self._dev_tasks -= 1
            if len(d.running_tasks) <= 1:
                for t in d.running_tasks.values():
                    if t.interfered:
                        relieved[t.task_id] = t.job_id
        if not reserved:
            self.finished_tasks.append(task.task_id)
        return relieved

    # -- execution ---------------------------------------------------------------------------------------
    def execute_job(self, job_id, delta_time, apply_interference: bool = False):
        """Start this node's share of a placed job. Returns (job or None, started task count); the
        job is returned once *all* of its tasks (on all nodes) run.

        ``apply_interference=False`` is parity mode: the reference tests a Task object against a
        set of task-id strings (infra/node.py:201), so its 1.2x slowdown never applies.
        Calibrated mode passes True and the co-located tasks really get longer."""
        job = self.placed_jobs.get(job_id)
        if job is None:
            raise ValueError(f"job {job_id} was not placed on node {self.node_id}")
        if self.node_id not in set(job.tasks_running_on.values()):
            raise ValueError(f"job {job_id} has no task on node {self.node_id}")
        crowded = set()
        for d in self.device_cache.values():
            if len(d.running_tasks) >= 2:
                crowded.update(d.running_tasks.keys())
        for task_id, where in job.tasks_running_on.items():
            if where != self.node_id:
                continue
            t = self.placed_tasks.pop(task_id)
            t.execute(delta_time)
            if apply_interference and t.task_id in crowded and not t.interfered:
                t.duration = t.duration * (1 + self.interference_factor)
                t.interfered = True
            self.running_tasks[task_id] = t
        ok, started = job.try_execute(delta_time)
        return (job if ok else None), started


class Rack:
    """A switch and the nodes under it."""

    def __init__(self, rack_id, bandwidth):
        self.rack_id = rack_id
        self.nodes: "OrderedDict[str, Node]" = OrderedDict()
        self.network_bandwidth = bandwidth

    def add_node(self, node: Node) -> None:
        if node.node_id in self.nodes:
            util.print_fn("Node already in rack", util.LOG_LEVEL_WARNING)
        else:
            self.nodes[node.node_id] = node


class Infrastructure:
    """The whole cluster, built from flags or a ``cluster_spec`` csv (header + one row with the
    five ``SPEC_KEYS`` columns; a spec lacking one of them yields an *empty* cluster, as in the
    reference, infra/infrastructure.py:91-92)."""

    def __init__(self, flags, rng=None):
        self.flags = flags
        self.rng = rng if rng is not None else global_rng()
        self.racks: "OrderedDict[str, Rack]" = OrderedDict()
        self.nodes: "OrderedDict[str, Node]" = OrderedDict()
        self.num_switch = flags.num_switch
        self.bandwidth = flags.bandwidth
        self.internode_latency = flags.internode_latency
        self.enable_network_costs = flags.enable_network_costs
        self.gpu_memory_capacity = flags.gpu_memory_capacity * 1024        # GiB -> MiB
        self.num_nodes_p_switch = flags.num_node_p_switch
        self.num_cpu_p_node = flags.num_cpu_p_node
        self.num_gpu_p_node = flags.num_gpu_p_node
        self.mem_p_node = flags.mem_p_node
        self.cluster_spec = flags.cluster_spec
        self.verbose = bool(getattr(flags, "print", False))
        self.racks_dist_map: Dict[str, List[Tuple[str, int]]] = {}
        self._setup_nodes()

    # -- construction --------------------------------------------------------------------------------
    def _setup_nodes(self) -> None:
        spec = self._resolve_spec_path(self.cluster_spec)
        if spec is None:
            self._init_nodes()
        else:
            self._init_from_spec_file(spec)

    @staticmethod
    def _resolve_spec_path(path: Optional[str]) -> Optional[str]:
        """The reference requires the path to exist relative to the cwd and then opens it relative
        to the repository root (infra/infrastructure.py:73-81); here: cwd first, then repo root."""
        if not path:
            return None
        if os.path.exists(path):
            return path
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        cand = os.path.join(root, path)
        return cand if os.path.exists(cand) else None

    def _init_from_spec_file(self, spec_file: str) -> None:
        assert "csv" in os.path.splitext(spec_file)[1]
        with open(spec_file, "r") as fh:
            reader = csv.DictReader(fh, delimiter=",")
            keys = reader.fieldnames or []
            if any(k not in keys for k in SPEC_KEYS):
                return
            for row in reader:
                self.num_switch = int(row["num_switch"])
                self.num_nodes_p_switch = int(row["num_node_p_switch"])
                self.num_gpu_p_node = int(row["num_gpu_p_node"])
                self.num_cpu_p_node = int(row["num_cpu_p_node"])
                self.mem_p_node = int(row["mem_p_node"])
        self._init_nodes()

    def _init_nodes(self) -> None:
        count = 0
        for rack_idx in range(self.num_switch):
            rack = Rack(str(rack_idx), self.bandwidth)
            for _ in range(self.num_nodes_p_switch):
                count += 1
                node = Node(rack.rack_id, str(count), self.gpu_memory_capacity, self.num_cpu_p_node,
                            self.num_gpu_p_node, self.mem_p_node, enable_pack=self.flags.pack,
                            rng=self.rng, verbose=self.verbose)
                self.nodes[node.node_id] = node
                rack.add_node(node)
            self.racks[rack.rack_id] = rack
        util.print_fn("num_racks in cluster: %d" % len(self.racks))
        if self.racks and self.nodes:
            first = next(iter(next(iter(self.racks.values())).nodes.values()))
            util.print_fn("num_node_p_rack in cluster: %d" % self.num_nodes_p_switch)
            util.print_fn("num_gpu_p_node in cluster: %d" % first.gpu_count)
            util.print_fn("num_cpu_p_node in cluster: %d" % first.cpu_count)
            util.print_fn("mem_p_node in cluster: %d" % first.mem_size)
        util.print_fn("Total nodes in cluster: %d " % len(self.nodes))
        util.print_fn("Total racks in cluster: %d " % len(self.racks))
        util.print_fn("--------------------------------- End of cluster spec ---------------------------------")

    # -- totals ------------------------------------------------------------------------------------------
    def get_available_cpu_count(self):
        return sum(n.cpu_free() for n in self.nodes.values())

    def get_available_mem_size(self):
        return sum(n.mem_free() for n in self.nodes.values())

    def get_total_gpus(self) -> int:
        return sum(n.gpu_count for n in self.nodes.values())

    def get_free_gpus(self) -> int:
        return sum(n.gpu_free() for n in self.nodes.values())

    def get_free_nodes(self) -> List[Node]:
        return [n for n in self.nodes.values() if n.is_free()]

    def get_racks_by_dist(self, other_rack_id) -> List[Tuple[str, int]]:
        """Racks ordered by ``|id - other|`` (stable), memoised per origin rack."""
        cached = self.racks_dist_map.get(other_rack_id)
        if cached:
            return cached
        dist = sorted(((r.rack_id, abs(int(r.rack_id) - int(other_rack_id))) for r in self.racks.values()),
                      key=lambda x: x[1])
        self.racks_dist_map[other_rack_id] = dist
        return dist

    def iter_devices(self) -> Iterable[Device]:
        for n in self.nodes.values():
            yield from n.device_cache.values()
