"""Counting cluster of the event-driven schedulers (reference: infra/cluster.py:19-148, :888-1483,
infra/switch.py:7-240; D11, D12). The reference's ``_Node`` class is undefined there
(infra/switch.py:27); it is reconstructed from the calls made on it.

Placement schemes (``--scheme``):
``count``    GPU counting only (``free_gpu``), no assignment.
``yarn``     MS-YARN: the first switch that can hold the job, all-or-nothing per switch; jobs larger
             than a node take ``floor(n/gpn)`` completely idle nodes plus one partial node.
``random`` / ``crandom``   random node per worker / consolidated random nodes (python ``random``).
``greedy``   most-free-GPU nodes first (the reference's body is a stub that only asserts).
``balance`` / ``cbalance`` minimise the maximum per-node network load (the LP of core/lp.py, solved
             here by :mod:`tiresias_b200.sim.legacy.lp`).
CPU accounting: 2 per worker, +4 per PS shard; a single-GPU job has no PS. Memory per multi-node
task: ``ps_mem + n*p_w_mem + worker_mem`` GB.
"""
from __future__ import annotations

import math
import random
from typing import Dict, List, Optional

from .jobs import JobRegistry, TJob

WORKER_CPU = 2
PS_CPU = 4


class TNode:
    def __init__(self, node_id: int, num_gpu: int, num_cpu: int, mem: float):
        self.id = node_id
        self.num_gpu, self.num_cpu, self.mem = num_gpu, num_cpu, mem
        self.init_node(num_gpu, num_cpu)

    def init_node(self, num_gpu: int = 0, num_cpu: int = 0, mem: Optional[float] = None) -> None:
        if num_gpu:
            self.num_gpu = num_gpu
        if num_cpu:
            self.num_cpu = num_cpu
        if mem is not None:
            self.mem = mem
        self.free_gpus = self.num_gpu
        self.free_cpus = self.num_cpu
        self.free_mem = self.mem
        self.network_in = 0.0
        self.network_out = 0.0

    def check_free_gpus(self) -> int:
        return self.free_gpus

    def check_free_cpus(self) -> int:
        return self.free_cpus

    def alloc_job_res(self, num_gpu: int, num_cpu: int) -> bool:
        if self.free_gpus < num_gpu or self.free_cpus < num_cpu:
            return False
        self.free_gpus -= num_gpu
        self.free_cpus -= num_cpu
        return True

    def alloc_cpus(self, num_cpu: int) -> bool:
        if self.free_cpus < num_cpu:
            return False
        self.free_cpus -= num_cpu
        return True

    def release_cpus(self, num_cpu: int) -> None:
        self.free_cpus = min(self.num_cpu, self.free_cpus + num_cpu)

    def release_gpus(self, num_gpu: int) -> bool:
        if self.free_gpus + num_gpu > self.num_gpu:
            return False
        self.free_gpus += num_gpu
        return True

    def release_job_gpu_cpu(self, num_gpu: int, num_cpu: int) -> None:
        self.free_gpus = min(self.num_gpu, self.free_gpus + num_gpu)
        self.free_cpus = min(self.num_cpu, self.free_cpus + num_cpu)

    def add_network_load(self, in_load: float = 0, out_load: float = 0) -> None:
        self.network_in = round(self.network_in + in_load, 1)
        self.network_out = round(self.network_out + out_load, 1)

    def release_network_load(self, in_load: float = 0, out_load: float = 0) -> None:
        self.network_in = max(0.0, round(self.network_in - in_load, 1))
        self.network_out = max(0.0, round(self.network_out - out_load, 1))

    def release_job_res(self, node_dict: dict) -> bool:
        self.release_job_gpu_cpu(node_dict["num_gpu"], node_dict["num_cpu"])
        self.free_mem = min(self.mem, self.free_mem + node_dict.get("mem", 0))
        net = node_dict.get("network", 0)
        self.release_network_load(net, net)
        return True


class TSwitch:
    def __init__(self, switch_id: int, num_node: int, num_gpu_p_node: int, num_cpu_p_node: int,
                 mem_p_node: float):
        self.id = switch_id
        self.num_node = num_node
        self.num_gpu_p_node, self.num_cpu_p_node, self.mem_p_node = num_gpu_p_node, num_cpu_p_node, mem_p_node
        self.node_list = [TNode(n, num_gpu_p_node, num_cpu_p_node, mem_p_node) for n in range(num_node)]

    # -- MS-YARN ---------------------------------------------------------------------------------------
    def try_single_node_alloc(self, jobs: JobRegistry, job: TJob) -> bool:
        need_gpu = job.num_gpu
        solo = not job.ps_network and need_gpu == 1
        need_cpu = need_gpu * (WORKER_CPU if solo else WORKER_CPU + PS_CPU)
        for node in self.node_list:
            if (node.free_gpus >= need_gpu and node.free_cpus >= need_cpu
                    and node.free_mem >= jobs.worker_mem and node.alloc_job_res(need_gpu, need_cpu)):
                node.free_mem -= jobs.worker_mem
                jobs.create_single_node_placement(job, self.id, node.id, need_gpu, need_cpu, jobs.worker_mem)
                return True
        return False

    def try_cross_node_alloc(self, jobs: JobRegistry, job: TJob) -> bool:
        gpn = self.num_gpu_p_node
        need = job.num_gpu
        full_nodes, tail_gpu = need // gpn, need % gpn
        per_gpu_cpu = WORKER_CPU + PS_CPU
        task_mem = jobs.ps_mem + need * jobs.p_w_mem + jobs.worker_mem
        chosen: List[TNode] = []
        for node in self.node_list:
            if len(chosen) == full_nodes:
                break
            if (node.free_gpus == node.num_gpu and node.free_cpus >= gpn * per_gpu_cpu
                    and node.free_mem >= task_mem * node.num_gpu):
                chosen.append(node)
        if len(chosen) < full_nodes:
            return False
        tail: Optional[TNode] = None
        if tail_gpu:
            for node in self.node_list:
                if node not in chosen and node.free_gpus >= tail_gpu \
                        and node.free_cpus >= tail_gpu * per_gpu_cpu and node.free_mem >= task_mem * tail_gpu:
                    tail = node
                    break
            if tail is None:
                return False
        model_size = job.model["total_size"]
        records = []
        shard = 0
        for node, local in [(n, n.num_gpu) for n in chosen] + ([(tail, tail_gpu)] if tail else []):
            node.alloc_job_res(local, local * per_gpu_cpu)
            node.free_mem -= task_mem * local
            traffic = round(model_size * local, 1)
            for _ in range(local):
                # a local PS shard talks to the (need - local) remote workers; local worker<->PS
                # exchanges never reach the NIC
                traffic += job.ps_network[shard] * (need - local)
                traffic -= job.ps_network[shard] * local
                traffic = round(traffic, 1)
                shard += 1
            node.add_network_load(traffic, traffic)
            records.append({"id": node.id, "num_gpu": local, "num_cpu": local * per_gpu_cpu,
                            "mem": task_mem * local, "network": traffic, "tasks": []})
        jobs.create_multi_nodes_placement(job, self.id, records)
        return True

    def ms_yarn_alloc_res(self, jobs: JobRegistry, job: TJob) -> bool:
        if job.num_gpu > self.num_gpu_p_node:
            return self.try_cross_node_alloc(jobs, job)
        return self.try_single_node_alloc(jobs, job)

    ms_yarn_alloc_gpus = ms_yarn_alloc_res

    def release_job_res(self, nodes: List[dict]) -> bool:
        for nd in nodes:
            if any(k not in nd for k in ("id", "num_gpu", "num_cpu", "tasks")):
                return False
            if not self.node_list[nd["id"]].release_job_res(nd):
                return False
        return True

    def release_gpus(self, nodes: List[dict]) -> bool:
        for nd in nodes:
            if "id" not in nd or "num_gpu" not in nd:
                return False
            if not self.node_list[nd["id"]].release_gpus(nd["num_gpu"]):
                return False
        return True


class TCluster:
    """The ``CLUSTER`` object."""

    GANDIVA_CLASSES = (1, 2, 4, 8, 12, 16, 24, 32, 64)

    def __init__(self, flags, jobs: JobRegistry, rng: Optional[random.Random] = None):
        self.flags = flags
        self.jobs = jobs
        self.rng = rng or random
        self.switch_list: List[TSwitch] = []
        self.gpu_list: List[float] = []
        self.free_nodes: List[TNode] = []
        self.node_g: Dict[int, List[dict]] = {k: [] for k in self.GANDIVA_CLASSES}
        self._buddy = None
        self.set_spec(flags.num_switch, flags.num_node_p_switch, flags.num_gpu_p_node,
                      flags.num_cpu_p_node, flags.mem_p_node)

    def set_spec(self, num_switch, num_node_p_switch, num_gpu_p_node, num_cpu_p_node, mem_p_node) -> None:
        self.num_switch, self.num_node_p_switch = num_switch, num_node_p_switch
        self.num_gpu_p_node, self.num_cpu_p_node, self.mem_p_node = num_gpu_p_node, num_cpu_p_node, mem_p_node
        self.num_node = num_switch * num_node_p_switch
        self.num_gpu = self.num_node * num_gpu_p_node
        self.num_cpu = self.num_node * num_cpu_p_node
        self.mem = self.num_node * mem_p_node
        self.free_gpu = self.num_gpu

    def init_infra(self) -> None:
        self.switch_list = [TSwitch(s, self.num_node_p_switch, self.num_gpu_p_node, self.num_cpu_p_node,
                                    self.mem_p_node) for s in range(self.num_switch)]
        if self.flags.schedule == "dlas-gpu-pack":
            self.init_dlas_pack_gpu()

    def print_cluster_spec(self) -> str:
        return ("#ofswitch: %d, #ofnode: %d, #ofgpu: %d, #ofcpu: %d, #ofmem: %d"
                % (self.num_switch, self.num_node, self.num_gpu, self.num_cpu, self.mem))

    def nodes(self):
        for s in self.switch_list:
            yield from s.node_list

    def empty_infra(self) -> None:
        self.free_gpu = self.num_gpu
        self._buddy = None
        for node in self.nodes():
            node.init_node(self.num_gpu_p_node, self.num_cpu_p_node, self.mem_p_node)
        if self.flags.schedule == "dlas-gpu-pack":
            self.init_dlas_pack_gpu()

    # -- fractional GPU packing (dlas-gpu-pack, D8) ---------------------------------------------------------
    def init_dlas_pack_gpu(self) -> None:
        self.gpu_list = [1] * self.num_gpu

    def free_gpu_util(self, job: TJob) -> bool:
        share = job.model["mem_util"]
        return sum(1 for g in self.gpu_list if g >= share) >= job.num_gpu

    def dlas_pack_get_gpu_util(self, job: TJob) -> bool:
        share = job.model["mem_util"]
        got = 0
        for i, g in enumerate(self.gpu_list):
            if g >= share:
                self.gpu_list[i] = g - share
                job.gpus.append(i)
                got += 1
            if job.num_gpu <= got:
                return True
        return False

    # -- placement dispatch ------------------------------------------------------------------------------------
    def try_get_job_res(self, job: TJob) -> bool:
        scheme = self.flags.scheme
        if scheme == "count":
            return self.none_placement(job)
        if scheme == "yarn":
            return self.ms_yarn_placement(job)
        if scheme == "random":
            return self.random_placement(job)
        if scheme == "crandom":
            return self.consolidate_random_placement(job)
        if scheme == "greedy":
            return self.greedy_placement(job)
        if scheme in ("balance", "cbalance"):
            from .lp import placement as lp_placement
            return lp_placement(self, job, consolidate=(scheme == "cbalance"))
        if scheme == "gandiva":
            return self.gandiva_placement(job)
        if scheme == "buddy":
            return self.buddy_placement(job)
        raise KeyError(f"unknown placement scheme {scheme!r}")

    def buddy_placement(self, job: TJob) -> bool:
        """One NVSwitch domain as the execution backend sees it (tiresias_b200.exec.buddy): a job runs on
        an aligned power-of-two block of GPUs, so a free GPU count is not enough — the block must exist.
        Lets the simulator reproduce the fragmentation the real 8xB200 replay experiences."""
        from ...exec.buddy import BuddyAllocator, normalize_gpus
        if self._buddy is None:
            self._buddy = BuddyAllocator(self.num_gpu)
        blk = self._buddy.alloc(normalize_gpus(job.num_gpu, self.num_gpu), str(job.job_id))
        if blk is None:
            return False
        job.placements.append({"buddy": blk})
        self.free_gpu -= blk[1]
        return True

    def none_placement(self, job: TJob) -> bool:
        if self.free_gpu >= job.num_gpu:
            self.free_gpu = int(self.free_gpu - job.num_gpu)
            return True
        return False

    def check_free_gpu(self) -> int:
        if self.flags.scheme == "count":
            return self.free_gpu
        return sum(n.free_gpus for n in self.nodes())

    def ms_yarn_placement(self, job: TJob) -> bool:
        return any(s.ms_yarn_alloc_res(self.jobs, job) for s in self.switch_list)

    def get_node_with_gid(self, gid: int) -> dict:
        switch = self.switch_list[gid // self.num_node_p_switch]
        return {"switch": switch, "node": switch.node_list[gid % self.num_node_p_switch]}

    # .. random / crandom .........................................................................................
    def _record_ps_worker_placements(self, job: TJob, w_nodes, w_switches, ps_nodes, ps_switches) -> None:
        """One placement record per worker and per PS shard, with the traffic each adds to its node
        (co-located worker/PS pairs exchange nothing over the network)."""
        num_w = job.num_gpu
        where: Dict[int, dict] = {}
        for n in w_nodes:
            where.setdefault(id(n), {"worker": 0, "ps": []})["worker"] += 1
        for i, n in enumerate(ps_nodes):
            where.setdefault(id(n), {"worker": 0, "ps": []})["ps"].append(i)
        for i in range(num_w):
            node = w_nodes[i]
            local = 0
            for j in where[id(node)]["ps"]:
                local = round(local + job.ps_network[j], 1)
            net = round(job.w_network[i] - local, 1)
            node.add_network_load(net, net)
            job.placements.append({"switch": w_switches[i], "nodes": [
                {"id": node.id, "num_gpu": 1, "num_cpu": WORKER_CPU, "network": net, "tasks": []}]})
        for i in range(len(ps_nodes)):
            node = ps_nodes[i]
            net = round(job.ps_network[i] * (num_w - where[id(node)]["worker"]), 1)
            node.add_network_load(net, net)
            job.placements.append({"switch": ps_switches[i], "nodes": [
                {"id": node.id, "num_gpu": 0, "num_cpu": PS_CPU, "network": net, "tasks": []}]})

    def random_placement(self, job: TJob) -> bool:
        num_ps, num_w = len(job.ps_network), job.num_gpu
        assert num_w == 1 or num_ps == num_w
        w_nodes, w_sw, ps_nodes, ps_sw = [], [], [], []

        def pick(ok, take):
            start = self.rng.randint(0, self.num_node - 1)
            for i in range(self.num_node):
                info = self.get_node_with_gid((start + i) % self.num_node)
                if ok(info["node"]):
                    take(info["node"])
                    return info
            return None

        def undo() -> bool:
            for n in w_nodes:
                n.release_job_gpu_cpu(1, WORKER_CPU)
            for n in ps_nodes:
                n.release_cpus(PS_CPU)
            return False

        for _ in range(num_w):
            info = pick(lambda n: n.free_gpus >= 1 and n.free_cpus >= WORKER_CPU,
                        lambda n: n.alloc_job_res(1, WORKER_CPU))
            if info is None:
                return undo()
            w_nodes.append(info["node"])
            w_sw.append(info["switch"].id)
        for _ in range(num_ps):
            info = pick(lambda n: n.free_cpus >= PS_CPU, lambda n: n.alloc_cpus(PS_CPU))
            if info is None:
                return undo()
            ps_nodes.append(info["node"])
            ps_sw.append(info["switch"].id)
        self._record_ps_worker_placements(job, w_nodes, w_sw, ps_nodes, ps_sw)
        return True

    def consolidate_random_placement(self, job: TJob) -> bool:
        """Whole nodes first (random switch, random node, wrap-around scan); whatever could not be
        consolidated is spread greedily; PS shards are shuffled over the chosen worker slots."""
        num_ps, num_w = len(job.ps_network), job.num_gpu
        assert num_w == 1 or num_ps == num_w
        gpn = self.num_gpu_p_node
        chunks = [gpn] * (num_w // gpn) + ([num_w % gpn] if num_w % gpn else [])
        cpu_per_gpu = WORKER_CPU if num_w == 1 else WORKER_CPU + PS_CPU
        slots: List[TNode] = []
        slot_sw: List[int] = []

        def scan(fn) -> bool:
            """Visit every node once starting at a random switch / node; stop when fn says so."""
            s0 = self.rng.randint(0, self.num_switch - 1)
            for ds in range(self.num_switch):
                sw = self.switch_list[(s0 + ds) % self.num_switch]
                n0 = self.rng.randint(0, self.num_node_p_switch - 1)
                for dn in range(self.num_node_p_switch):
                    if fn(sw, sw.node_list[(n0 + dn) % self.num_node_p_switch]):
                        return True
            return False

        done = 0
        for need in chunks:
            def whole(sw, node, need=need) -> bool:
                if node.free_gpus >= need and node.free_cpus >= need * cpu_per_gpu:
                    node.alloc_job_res(need, need * cpu_per_gpu)
                    slots.extend([node] * need)
                    slot_sw.extend([sw.id] * need)
                    return True
                return False
            if not scan(whole):
                break
            done += 1
        remain = sum(chunks[done:])
        if remain:
            def undo() -> bool:
                for n in slots:
                    n.release_job_gpu_cpu(1, cpu_per_gpu)
                return False
            if remain == 1:
                return undo()
            left = [remain]

            def spread(sw, node) -> bool:
                if node.free_gpus >= 1 and node.free_cpus >= cpu_per_gpu:
                    take = min(node.free_gpus, node.free_cpus // cpu_per_gpu, left[0])
                    node.alloc_job_res(take, take * cpu_per_gpu)
                    slots.extend([node] * take)
                    slot_sw.extend([sw.id] * take)
                    left[0] -= take
                return left[0] == 0
            scan(spread)
            if left[0]:
                return undo()
        order = list(range(num_ps))
        self.rng.shuffle(order)
        ps_nodes = [slots[i] for i in order]
        ps_sw = [slot_sw[i] for i in order]
        self._record_ps_worker_placements(job, slots, slot_sw, ps_nodes, ps_sw)
        # the CPU share of the PS shards was reserved together with the workers above
        for rec in job.placements[-num_ps:] if num_ps else []:
            rec["nodes"][0]["num_cpu"] = PS_CPU
        return True

    def greedy_placement(self, job: TJob) -> bool:
        """Fill the nodes with the most free GPUs first (fewest nodes => least cross-node traffic)."""
        num_w = job.num_gpu
        cpu_per_gpu = WORKER_CPU if num_w == 1 else WORKER_CPU + PS_CPU
        order = sorted(((sw, n) for sw in self.switch_list for n in sw.node_list),
                       key=lambda p: -min(p[1].free_gpus, p[1].free_cpus // cpu_per_gpu))
        plan, left = [], num_w
        for sw, node in order:
            take = min(node.free_gpus, node.free_cpus // cpu_per_gpu, left)
            if take > 0:
                plan.append((sw, node, take))
                left -= take
            if left == 0:
                break
        if left:
            return False
        shard = 0
        for sw, node, take in plan:
            node.alloc_job_res(take, take * cpu_per_gpu)
            traffic = 0.0
            if num_w > 1 and len(plan) > 1:
                traffic = round(job.model["total_size"] * take, 1)
                for _ in range(take):
                    traffic = round(traffic + job.ps_network[shard] * (num_w - take)
                                    - job.ps_network[shard] * take, 1)
                    shard += 1
            node.add_network_load(traffic, traffic)
            job.placements.append({"switch": sw.id, "nodes": [
                {"id": node.id, "num_gpu": take, "num_cpu": take * cpu_per_gpu, "network": traffic, "tasks": []}]})
        return True

    # -- release -------------------------------------------------------------------------------------------------
    def release_job_res(self, job: TJob) -> bool:
        if self.flags.schedule == "dlas-gpu-pack":
            share = job.model["mem_util"]
            for idx in job.gpus:
                self.gpu_list[idx] = min(1, self.gpu_list[idx] + share)
            del job.gpus[:]
            job.status = "END"
            return True
        if self.flags.scheme == "count":
            self.free_gpu = min(self.num_gpu, self.free_gpu + job.num_gpu)
            job.status = "END"
            return True
        if self.flags.scheme == "buddy":
            for placement in job.placements:
                if self._buddy is not None and "buddy" in placement:
                    self._buddy.free(tuple(placement["buddy"]))
                    self.free_gpu = min(self.num_gpu, self.free_gpu + placement["buddy"][1])
            del job.placements[:]
            job.status = "END"
            return True
        for placement in job.placements:
            if "switch" not in placement or "nodes" not in placement:
                job.status = "ERROR"
                return False
            if not self.switch_list[placement["switch"]].release_job_res(placement["nodes"]):
                job.status = "ERROR"
                return False
        job.status = "END"
        return True

    # -- legacy Gandiva node sets (D9) ---------------------------------------------------------------------------------
    def init_gandiva_nodes(self) -> None:
        self.free_nodes = list(self.nodes())
        assert len(self.free_nodes) == self.num_node

    def _new_node_set(self, num_gpu: int) -> Optional[dict]:
        num_node = int(math.ceil(num_gpu / self.num_gpu_p_node))
        if num_node > len(self.free_nodes):
            return None
        return {"nodes": [self.free_nodes.pop(0) for _ in range(num_node)], "gpu_u": [], "jobs": [],
                "concurrency": 0, "capacity": int(num_node * self.num_gpu_p_node / num_gpu),
                "num_gpus": num_gpu, "num_jobs": 0, "util": 0}

    @staticmethod
    def _ns_add(node_set: dict, job: TJob) -> None:
        node_set["util"] = round(node_set["util"] + round(job.model["mem_util"], 2), 2)
        node_set["jobs"].append(job)
        node_set["num_jobs"] += 1

    def gandiva_placement(self, job: TJob) -> bool:
        """Least-loaded node set of the job's GPU class if it has room; else a fresh node set from
        the free nodes; else over-subscribe the least-loaded set (time slicing); else pend."""
        if job.num_gpu not in self.node_g:
            raise ValueError("job[%d] needs %d GPUs: no gandiva class" % (job.job_idx, job.num_gpu))
        sets = self.node_g[job.num_gpu]
        share = round(job.model["mem_util"], 2)
        if sets:
            if round(sets[0]["util"] + share, 2) < sets[0]["capacity"]:
                self._ns_add(sets[0], job)
                sets.sort(key=lambda ns: ns["util"])
                return True
            sets.append(sets.pop(0))              # a full head set goes to the back, unsorted
        fresh = self._new_node_set(job.num_gpu)
        if fresh is not None:
            self._ns_add(fresh, job)
            sets.append(fresh)
            sets.sort(key=lambda ns: ns["util"])
            return True
        if sets:
            self._ns_add(sets[0], job)
            sets.sort(key=lambda ns: ns["util"])
            return True
        return False

    def _rebalance(self, sets: List[dict], moved: List[TJob]) -> None:
        for job in moved:
            self._ns_add(sets[0], job)
            sets.sort(key=lambda ns: ns["util"])

    def gandiva_node_set_shrink(self, num_gpu: int, sets: List[dict], count: int) -> int:
        count = min(count, len(sets) - 1)
        moved: List[TJob] = []
        for _ in range(max(0, count)):
            ns = sets.pop(0)
            moved.extend(ns["jobs"])
            self.free_nodes.extend(ns["nodes"])
        if count > 0:
            self._rebalance(sets, moved)
        return max(0, count)

    def gandiva_node_set_expand(self, num_gpu: int, sets: List[dict], count: int) -> int:
        made = 0
        for _ in range(count):
            fresh = self._new_node_set(num_gpu)
            if fresh is None:
                break
            sets.append(fresh)
            made += 1
        if made:
            moved: List[TJob] = []
            for ns in sets:
                moved.extend(ns["jobs"])
                ns["jobs"], ns["util"], ns["num_jobs"] = [], 0, 0
            self._rebalance(sets, moved)
        return made

    def gandiva_node_set_adjust(self) -> None:
        """Grow / shrink each GPU class's node sets towards its share of the total GPU demand."""
        demand = {g: sum(len(ns["jobs"]) for ns in sets) * g for g, sets in self.node_g.items()}
        occupied = {g: sum(len(ns["nodes"]) for ns in sets) * self.num_gpu_p_node
                    for g, sets in self.node_g.items()}
        total = sum(demand.values())
        if total == 0:
            return
        for g, sets in self.node_g.items():
            if demand[g] == 0:
                continue
            plan = int(math.floor(demand[g] / total * self.num_gpu))
            diff = min(plan, demand[g]) - occupied[g]
            if diff > 0:
                self.gandiva_node_set_expand(g, sets, int(math.ceil(diff / g)))
            elif diff < 0:
                self.gandiva_node_set_shrink(g, sets, int(math.ceil(-diff / g)))

    def time_slicing_execute(self, cur_time: int, jobs: JobRegistry, log, time_diff: int) -> bool:
        """Advance the jobs at the head of every node set by ``time_diff`` (as many as fit the set's
        capacity run concurrently), rotate the heads every 60 s, retire finished jobs and free empty
        node sets. Returns whether nodes were released."""
        released = False
        rotate = cur_time % 60 == 0
        used_gpus = 0
        for g, sets in self.node_g.items():
            empty = []
            for ns in sets:
                concurrency, load = 0, 0
                for j in ns["jobs"]:
                    load += j.model["mem_util"]
                    if load > ns["capacity"]:
                        break
                    concurrency += 1
                ns_used = g if len(ns["jobs"]) * g > self.num_gpu_p_node else len(ns["jobs"]) * g
                used_gpus += ns_used
                finished = []
                for i, j in enumerate(ns["jobs"]):
                    if i >= max(concurrency, 1):
                        break
                    j.executed_time += time_diff
                    j.total_executed_time = j.executed_time
                    if j.executed_time >= j.duration:
                        j.status = "END"
                        finished.append(j)
                if rotate and len(ns["jobs"]) > concurrency:
                    turn = max(concurrency, 1)
                    ns["jobs"] = ns["jobs"][turn:] + ns["jobs"][:turn]
                for j in finished:
                    jobs.running_jobs.remove(j)
                    jobs.completed_jobs.append(j)
                    ns["jobs"].remove(j)
                    ns["num_jobs"] -= 1
                    log.job_complete(j, cur_time)
                if not ns["jobs"]:
                    self.free_nodes.extend(ns["nodes"])
                    empty.append(ns)
                    used_gpus -= ns_used
                    released = True
            for ns in empty:
                sets.remove(ns)
        free = len(self.free_nodes)
        log.gandiva_checkpoint(jobs, cur_time, free, used_gpus,
                               int(self.num_gpu - used_gpus - free * self.num_gpu_p_node),
                               len(jobs.pending_jobs), len(jobs.running_jobs),
                               *[len(self.node_g[g]) for g in (1, 2, 4, 8, 16, 32, 64)])
        return released
