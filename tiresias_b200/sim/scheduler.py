"""The tick loop of the live simulator (reference: core/scheduling/schedule.py:12-213, C4).

One tick, in the reference's order (the order decides ``cluster.csv``):
 1. arrivals due by ``delta`` enter the queues (durations scaled by 0.5),
 2. if anything is queued and a node is free: one scheduling attempt, at most one job starts,
 3. clock advances; queued jobs age, running tasks progress, horus+ credits refresh,
 4. finished jobs release their resources,
 5. optional migration scan (after tick 100), policy plug-in (gandiva time slicing),
 6. one ``cluster.csv`` row.
The loop ends when no arrivals remain and nothing runs — jobs still queued then are abandoned (Q3).

New here: every finished job also yields a ``jobs.csv`` row (the reference's live path logs no
per-job data, so average JCT could not be computed from its output), and ``summary()`` reports
makespan / average JCT / average pending time for BASELINE.json's headline table.
"""
from __future__ import annotations

import time
from typing import Dict, Optional

from . import util
from .infra import sample_utilizations
from .logs import ArrayLike, LogInfo
from .network import calculate_network_costs
from .placement import placement_algorithms
from .policies import plugin_algorithms, scheduling_algorithms

SCALE_FACTOR = 0.5             # core/scheduling/schedule.py:187
MIGRATION_WARMUP = 100         # core/scheduling/schedule.py:196


class Scheduler:
    def __init__(self, infrastructure, jobs_manager, log_manager, enable_migration: bool = False,
                 calibration: Optional[Dict] = None):
        self.infrastructure = infrastructure
        self.jobs_manager = jobs_manager
        self.log_manager = log_manager
        self.placement = infrastructure.flags.scheme
        self.schedule = infrastructure.flags.schedule
        self.enable_migration = enable_migration
        self.calibration = calibration
        self.parity = getattr(infrastructure.flags, "mode", "parity") != "calibrated"
        self.agent = None                      # hook for sim.env.Environment
        self.delta_time = 0
        self.steps = 0
        self.last_info: Optional[LogInfo] = None
        self._remaining = 0
        self._queuing = 0
        self._running = 0
        self.wall_seconds = 0.0
        if self.schedule not in scheduling_algorithms:
            raise KeyError(f"schedule {self.schedule!r} is not a tick policy "
                           f"(have {sorted(scheduling_algorithms)})")
        if self.placement not in placement_algorithms:
            raise KeyError(f"scheme {self.placement!r} is not a tick placement "
                           f"(have {sorted(placement_algorithms)})")

    # -- small accessors kept from the reference ------------------------------------------------------
    def add_rack(self, rack) -> None:
        self.infrastructure.racks[rack.rack_id] = rack

    def collate_all_nodes(self):
        return self.infrastructure.nodes

    def num_free_nodes(self) -> int:
        return sum(1 for n in self.infrastructure.nodes.values() if n.is_free())

    def unfinished_node_count(self) -> int:
        return sum(1 for n in self.infrastructure.nodes.values() if n.running_tasks)

    # -- one scheduling attempt -------------------------------------------------------------------------
    def _placement_scheme_name(self) -> str:
        """The reference hands the *schedule* name to the placement as its scoring scheme (Q7). In
        parity mode that is kept; when the schedule has no scoring function of its own (e.g.
        ``--scheme horus --schedule fifo``, a KeyError in the reference) the placement's own name is
        used so the combination works."""
        from .scoring import score_fn
        return self.schedule if self.schedule in score_fn else self.placement

    def _schedule(self, delta) -> None:
        if self.num_free_nodes() < 1:
            return
        jm = self.jobs_manager
        jobs_all = jm.total_jobs(delta)
        policy = scheduling_algorithms[self.schedule]
        placement = placement_algorithms[self.placement]
        nodes, job, success = policy(self._placement_scheme_name(), placement, self.infrastructure, jm, delta)
        if success:
            if self.infrastructure.enable_network_costs:
                extras = calculate_network_costs(self.infrastructure, job, self.calibration)
                original = job.duration
                job.add_network_costs(extras)
                util.print_fn("Job %s : Original duration %f , New duration %f"
                              % (job.job_id, original, job.duration))
            self.add_to_running(nodes, job.job_id, delta)
            assert jobs_all - 1 == jm.total_jobs(delta), (
                "Expected %d - Got %d" % (jobs_all - 1, jm.total_jobs(delta)))
        else:
            assert jobs_all == jm.total_jobs(delta)

    def add_to_running(self, nodes, job_id, delta_time) -> None:
        for node_id, node in nodes.items():
            self.jobs_manager.start_job(node, job_id, delta_time)
            assert node_id in self.jobs_manager.busy_nodes

    # -- completion -----------------------------------------------------------------------------------------
    def release_finished_jobs(self, current_time) -> None:
        jm = self.jobs_manager
        nodes = self.infrastructure.nodes
        for job in jm.prepare_finish_tasks(current_time):
            done = False
            for task_id, node_id in job.tasks_running_on.items():
                node = nodes[node_id]
                task = node.running_tasks.pop(task_id)
                assert not task.finished
                job.task_finished(task_id)
                relieved = node.release_allocated_resources(task)
                if relieved:
                    jm.reset_interference(relieved)
                done = job.try_finished()
                if done:
                    jm.running_jobs.pop(job.job_id)
                    if job.job_id not in jm.finished_jobs:
                        jm.finished_jobs[job.job_id] = job
                        job.end_time = current_time
                        self._log_job(job, current_time)
                if not node.running_tasks:
                    jm.busy_nodes.remove(node_id)
            assert done

    def _log_job(self, job, now) -> None:
        start = job.first_start_time if job.first_start_time is not None else job.start_time
        row = [now, job.job_id, int(job.gpus), job.submit_time, start, now, now - start,
               now - job.submit_time, job.duration, job.total_pending, job.preempt_count,
               job.promote_count]
        self.log_manager.job_complete(row)

    # -- migration ---------------------------------------------------------------------------------------------
    def _scan_for_migrate(self) -> None:
        """Find an over-packed device (>2 tasks) while idle devices exist and requeue its most
        demanding job. Parity mode reproduces the reference, where the overload counter is never
        updated so nothing is ever migrated; calibrated mode performs the migration."""
        idle_devices = []
        worst, worst_load = None, 0
        for n in self.infrastructure.nodes.values():
            if self.parity and not n.is_idle():
                continue
            for d in n.device_cache.values():
                if d.is_idle():
                    idle_devices.append((n.node_id, d.device_id))
                else:
                    load = len(d.running_tasks)
                    if load > worst_load:
                        worst = d
                        if not self.parity:
                            worst_load = load
        if not idle_devices or worst is None or worst_load <= 2:
            return
        target, target_util = None, 0
        for t in worst.running_tasks.values():
            if t.gpu_utilization_avg > target_util:
                target, target_util = t.job_id, t.gpu_utilization_avg
        if target is not None and target in self.jobs_manager.running_jobs:
            util.logger.info("*******preempting: %s on device %s  " % (target, worst.device_id))
            self.jobs_manager.preempt(target, self.infrastructure)

    # -- logging ---------------------------------------------------------------------------------------------------
    def construct_info(self) -> LogInfo:
        nodes = self.infrastructure.nodes
        cached = getattr(self, "_device_list", None)
        if cached is None or cached[0] != len(nodes):
            # devices never change after the cluster is built: flatten once (node order, device order)
            devices = [d for n in nodes.values() for d in n.device_cache.values()]
            cached = self._device_list = (len(nodes), devices, sum(d.memory for d in devices))
        _, devices, mem_cap = cached
        idle_nodes = 0
        for n in nodes.values():
            if not (n.running_tasks or n.placed_tasks or n.placed_jobs):        # Node.is_idle()
                idle_nodes += 1
        busy_nodes = len(nodes) - idle_nodes
        busy = [d for d in devices if d.running_tasks]
        busy_gpus = len(busy)
        idle_gpus = len(devices) - busy_gpus
        util_sum, util_is_array = 0.0, False
        mem_sum = 0
        # one generator call for the whole cluster (same stream order as device-by-device draws)
        for d, (u, drawn) in zip(busy, sample_utilizations(busy, self.infrastructure.rng)):
            util_sum += u
            util_is_array = util_is_array or drawn
            mem_sum += d.get_current_memory()
        avg_util = util_sum / (idle_gpus + busy_gpus)
        avg_util = ArrayLike(avg_util) if util_is_array else avg_util
        avg_mem = mem_sum / mem_cap
        jm = self.jobs_manager
        avg_p, med_p, max_p = jm.pending_time_infos()
        return LogInfo(idle_nodes, busy_nodes, busy_gpus, idle_gpus, avg_util, avg_mem, avg_p, med_p,
                       max_p, len(jm.running_jobs), jm.queuing_jobs(), len(jm.finished_jobs))

    _construct_info = construct_info

    # -- the loop -------------------------------------------------------------------------------------------------------
    def prime(self) -> None:
        jm = self.jobs_manager
        self.delta_time = 0
        self._remaining = jm.remaining_jobs(0)
        self._queuing = jm.queuing_jobs(0)
        self._running = len(jm.running_jobs)
        self.steps = 0

    def has_work(self) -> bool:
        return self._remaining + self._running > 0

    def tick(self, action=None) -> bool:
        """Run one tick; returns whether the loop would continue."""
        jm = self.jobs_manager
        delta = self.delta_time
        jm.gen_jobs(delta, scale_factor=SCALE_FACTOR)
        if jm.queuing_jobs(delta) > 0:
            self._schedule(delta)
        self._remaining = jm.remaining_jobs(delta)
        self._queuing = jm.queuing_jobs(delta)
        self.delta_time = delta = delta + 1
        jm.step()
        self.release_finished_jobs(delta)
        self._running = len(jm.running_jobs)
        if self.enable_migration and delta > MIGRATION_WARMUP:
            self._scan_for_migrate()
        plugin = plugin_algorithms.get(self.schedule)
        if plugin is not None:
            plugin(self.infrastructure, jm)
        self.steps += 1
        self.last_info = self.construct_info()
        self.log_manager.step_cluster(self.last_info, delta)
        if self.infrastructure.verbose:
            util.print_fn("Remaining jobs: %d, Queuing Jobs: %d Running Jobs: %d Finished Jobs %d"
                          % (self._remaining, self._queuing, self._running, len(jm.finished_jobs)))
        return self.has_work()

    def start(self) -> None:
        t0 = time.time()
        self.prime()
        while self.has_work():
            self.tick()
        self.wall_seconds = time.time() - t0
        s = self.summary()
        util.print_fn("Remaining jobs: %d, Queuing Jobs: %d Running Jobs: %d Finished Jobs %d"
                      % (self._remaining, self._queuing, self._running, s["finished_jobs"]))
        util.print_fn("Makespan: %d ticks, avg JCT %.2f, avg pending %.2f"
                      % (s["makespan"], s["avg_jct"], s["avg_pending"]))
        util.print_fn("Total Time Taken in seconds: %d" % self.wall_seconds)

    def summary(self) -> Dict[str, float]:
        done = list(self.jobs_manager.finished_jobs.values())
        n = len(done)
        jct = sum(j.end_time - j.submit_time for j in done) / n if n else float("nan")
        pend = sum(j.total_pending for j in done) / n if n else float("nan")
        return {"makespan": self.delta_time, "finished_jobs": n, "avg_jct": jct, "avg_pending": pend,
                "queued_at_end": self.jobs_manager.queuing_jobs(), "wall_seconds": self.wall_seconds,
                "ticks": self.steps}
