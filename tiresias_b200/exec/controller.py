"""Real-time scheduler of the execution backend.

Replays a README-schema trace (``job_id, num_gpu, submit_time, iterations, model_name, duration``) against
the worker pool: jobs arrive at ``submit_time * time_scale`` seconds of wall clock, a policy decides what
runs on which GPU block, finished / preempted jobs report back through the store. Policies mirror the
simulator's (tiresias_b200.sim.legacy.schedulers):

``fifo``      non-preemptive; waiting jobs are tried in arrival order until the first that does not fit
``sjf``       preemptive, fewest GPUs first
``shortest``  preemptive, least remaining iterations first
``dlas-gpu``  preemptive two-level (or more) MLFQ on attained GPU-seconds with ``queue_limit`` thresholds
``gittins``   preemptive, highest Gittins index of the attained GPU-time first
``gandiva``   fifo with Gandiva's time slicing (reference core/scheduling/algorithm.py:418-438): while jobs are
              waiting, a job that has held its GPUs for a whole quantum is preempted and re-queued. The reference
              re-inserts it at the FRONT of its list queue, where its fifo would hand the GPUs
              straight back to it; the executor re-queues at the BACK, i.e. real round-robin slicing

``pack=True`` executes what the reference's device model allows (infra/device.py:49-59: up to 3 tasks per
device): single-GPU jobs may share a GPU with up to ``max_share - 1`` other single-GPU jobs — they run as
threads of that GPU's worker process, same CUDA context, own streams, really concurrent. A free GPU is
preferred; only when none is left does a job join the shared GPU with the fewest tenants. Multi-GPU jobs
keep exclusive aligned blocks.

Preemption is executable here (the reference only flips a flag, core/jobs/jobs_manager.py:150-173): the
job's ranks snapshot their PS shards and the job later resumes from the iteration it reached, possibly on
other GPUs.
"""
from __future__ import annotations

import json
import time
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

from .buddy import Block, BuddyAllocator, normalize_gpus


def _mean(values) -> Optional[float]:
    vals = [v for v in values if v is not None]
    return sum(vals) / len(vals) if vals else None


@dataclass
class ExecJob:
    job_id: str
    model: str
    gpus: int
    iters: int
    submit: float                      # seconds after replay start
    batch: int
    status: str = "WAITING"            # WAITING | RUNNING | PREEMPTING | DONE | FAILED
    done: int = 0
    attempt: int = 0
    block: Optional[Block] = None
    first_start: Optional[float] = None
    last_start: Optional[float] = None
    end: Optional[float] = None
    attained: float = 0.0              # seconds of service received (all attempts)
    preempt: int = 0
    q_id: int = 0
    rank: float = 0.0                  # gittins index of the attained service (policy "gittins")
    results: List[Dict] = field(default_factory=list)

    def jct(self) -> Optional[float]:
        return None if self.end is None else self.end - self.submit


class Controller:
    def __init__(self, store, num_gpus: int, rows: Sequence[Dict], policy: str = "fifo", comm: str = "ps",
                 batch: int = 32, time_scale: float = 1.0, ckpt_dir: str = "/dev/shm/tb200_ckpt",
                 queue_limit: Sequence[float] = (20.0,), check_every: int = 10, poll_s: float = 0.02,
                 job_timeout_s: float = 600.0, step_sleep: float = 0.0, graph: bool = True, log=print,
                 job_dist_s: Optional[Sequence[float]] = None, gittins_delta_s: float = 2.0,
                 est_iter_s: float = 0.05, iter_table: Optional[Dict] = None, pack: bool = False,
                 max_share: int = 3, quantum_s: float = 10.0, interference: Optional[Sequence[float]] = None):
        self.store, self.num_gpus, self.policy, self.comm = store, num_gpus, policy, comm
        self.time_scale, self.ckpt_dir, self.check_every, self.poll_s = time_scale, ckpt_dir, check_every, poll_s
        self.queue_limit = list(queue_limit)
        self.job_timeout_s, self.step_sleep, self.graph = job_timeout_s, step_sleep, graph
        self.log = log
        self.alloc = BuddyAllocator(num_gpus)
        self.seq = [0] * num_gpus
        self.jobs: List[ExecJob] = []
        for r in sorted(rows, key=lambda x: (x["submit_time"], x["job_id"])):
            self.jobs.append(ExecJob(str(r["job_id"]), r.get("model_name") or "resnet50",
                                     normalize_gpus(r["num_gpu"], num_gpus), max(1, int(r["iterations"])),
                                     float(r["submit_time"]) * time_scale, int(r.get("batch") or batch)))
        # gittins: the index table over GPU-time samples in milliseconds (sim.legacy.gittins, PyTorch). Without
        # a distribution the trace's own jobs are the samples: iterations x estimated iteration time x GPUs —
        # what the simulator falls back to as well (sim/legacy/engine.py).
        # iteration-time estimates per (model, GPUs): from a calibration when one is given, refined by every
        # finished attempt; what "shortest" ranks remaining work with and what gittins builds its samples from
        self.est_iter_s = est_iter_s
        self.iter_table: Dict[str, Dict[str, float]] = {m: dict(v) for m, v in (iter_table or {}).items()}
        self.gittins_table = None
        self.gittins_delta_s = max(1e-3, float(gittins_delta_s))
        self._next_unit = self.gittins_delta_s
        self.gittins_samples_ms: Optional[List[int]] = None
        if policy == "gittins":
            from ..sim.legacy import gittins as gi
            samples = list(job_dist_s) if job_dist_s else [j.iters * self._iter_s(j) * j.gpus for j in self.jobs]
            self.gittins_samples_ms = [max(1, int(round(v * 1000.0))) for v in samples]
            self.gittins_table = gi.gittins_table(self.gittins_samples_ms,
                                                  max(1, int(round(gittins_delta_s * 1000.0))))
        self.pack, self.max_share, self.quantum_s = bool(pack), max(1, int(max_share)), float(quantum_s)
        self.interference = list(interference) if interference else None   # loopback: emulated slowdown per tenancy
        self.tenants: Dict[int, List[str]] = {}    # GPU -> ids of the single-GPU jobs sharing it (pack mode)
        self.t0 = 0.0
        self.pending: List[ExecJob] = []           # fifo: arrived, could not start yet
        self.queues: List[List[ExecJob]] = []      # dlas-gpu: the MLFQ, as explicit lists like the simulator's
        self.seen: set = set()

    def _iter_s(self, job: ExecJob) -> float:
        per = self.iter_table.get(job.model)
        if not per:
            return self.est_iter_s
        keys = sorted(int(k) for k in per)
        pick = max([k for k in keys if k <= job.gpus], default=keys[0])
        return float(per[str(pick)])

    # -- pool plumbing -------------------------------------------------------------------------------
    def wait_ready(self, timeout_s: float = 600.0) -> None:
        self.store.wait([f"ready/{r}" for r in range(self.num_gpus)], __import__("datetime").timedelta(seconds=timeout_s))

    def _send(self, rank: int, cmd: Dict) -> None:
        self.store.set(f"cmd/{rank}/{self.seq[rank]}", json.dumps(cmd))
        self.seq[rank] += 1

    def shutdown(self) -> None:
        for r in range(self.num_gpus):
            self._send(r, {"op": "exit"})

    def _start(self, job: ExecJob, now: float) -> bool:
        shared = self.pack and job.gpus == 1
        blk = self.alloc.alloc(job.gpus, job.job_id)
        if blk is None and shared:
            # no free GPU: join the shared GPU with the fewest tenants, if one has room
            room = [(len(t), g) for g, t in self.tenants.items() if 0 < len(t) < self.max_share]
            if room:
                blk = (min(room)[1], 1)
        if blk is None:
            return False
        if shared:
            self.tenants.setdefault(blk[0], []).append(job.job_id)
        job.block, job.status, job.last_start = blk, "RUNNING", now
        if job.first_start is None:
            job.first_start = now
        cmd = {"op": "run", "job": job.job_id, "attempt": job.attempt, "model": job.model, "batch": job.batch,
               "iters": job.iters, "done": job.done, "mode": self.comm, "block": list(blk),
               "resume": job.attempt > 0, "ckpt_dir": self.ckpt_dir, "check_every": self.check_every,
               "step_sleep": self.step_sleep, "graph": self.graph, "shared": shared,
               "interference": self.interference}
        for r in range(blk[0], blk[0] + blk[1]):
            self._send(r, cmd)
        self.log(f"[{now:8.2f}s] start job {job.job_id} ({job.model}, {job.gpus} GPU, it {job.done}/{job.iters}) "
                 f"on GPUs {blk[0]}..{blk[0] + blk[1] - 1}" + (" (resume)" if job.attempt else "")
                 + (f" (shares the GPU with {len(self.tenants[blk[0]]) - 1})" if shared and len(self.tenants[blk[0]]) > 1 else ""))
        return True

    def _release(self, job: ExecJob) -> None:
        """Give the job's GPUs back; a shared GPU returns to the allocator when its last tenant leaves."""
        if self.pack and job.gpus == 1 and job.block is not None and job.job_id in self.tenants.get(job.block[0], ()):
            left = self.tenants[job.block[0]]
            left.remove(job.job_id)
            if not left:
                self.alloc.free(job.block)
        elif job.block is not None:
            self.alloc.free(job.block)

    def _request_preempt(self, job: ExecJob, now: float) -> None:
        if job.status == "RUNNING":
            self.store.set(f"preempt/{job.job_id}/{job.attempt}", "1")
            job.status = "PREEMPTING"
            self.log(f"[{now:8.2f}s] preempt job {job.job_id}")

    def _collect(self, now: float) -> bool:
        """Fold in results of attempts whose ranks have all reported. Returns True if anything changed."""
        changed = False
        for job in self.jobs:
            if job.status not in ("RUNNING", "PREEMPTING"):
                continue
            keys = [f"res/{job.job_id}/{job.attempt}/{r}" for r in range(job.gpus)]
            if not self.store.check(keys):
                if now - job.last_start > self.job_timeout_s:           # watchdog
                    job.status, job.end = "FAILED", now
                    self._release(job)
                    changed = True
                    self.log(f"[{now:8.2f}s] job {job.job_id} timed out")
                continue
            res = [json.loads(self.store.get(k)) for k in keys]
            job.results.append({"attempt": job.attempt, "block": list(job.block), "ranks": res,
                                "t_start": job.last_start, "t_end": now})
            self._release(job)
            job.attained += now - job.last_start
            job.block = None
            changed = True
            if any(r["status"] == "failed" for r in res):
                job.status, job.end = "FAILED", now
                self.log(f"[{now:8.2f}s] job {job.job_id} FAILED: {[r.get('error') for r in res if r['status'] == 'failed'][:1]}")
            elif all(r["status"] == "done" for r in res):
                job.status, job.end, job.done = "DONE", now, job.iters
                if res[0].get("iter_ms") and str(job.gpus) not in self.iter_table.get(job.model, {}):
                    self.iter_table.setdefault(job.model, {})[str(job.gpus)] = res[0]["iter_ms"] / 1e3
                self.log(f"[{now:8.2f}s] job {job.job_id} done (JCT {job.jct():.2f}s, "
                         f"{(res[0].get('iter_ms') or 0.0):.2f} ms/iter)")
            else:
                job.done = min(r["done"] for r in res)
                job.status = "WAITING"
                job.attempt += 1
                job.preempt += 1
        return changed

    # -- policies -------------------------------------------------------------------------------------------
    def _schedule(self, now: float) -> None:
        arrived = [j for j in self.jobs if j.submit <= now]
        waiting = [j for j in arrived if j.status == "WAITING"]
        running = [j for j in arrived if j.status in ("RUNNING", "PREEMPTING")]
        if self.policy == "fifo":
            # the simulator's sim_job_events (reference run_sim.py:1533-1599), statement for statement:
            # pending jobs first, in list order, until one does not fit (removing the started job from the
            # list being iterated skips its successor for this event, as there); then the new arrivals,
            # which start at once if they fit and pend otherwise; finally pending is sorted by num_gpu.
            for j in self.pending:
                if self._start(j, now):
                    self.pending.remove(j)
                else:
                    break
            for j in waiting:
                if j.job_id in self.seen:
                    continue
                self.seen.add(j.job_id)
                if j.status == "WAITING" and not self._start(j, now):
                    self.pending.append(j)
            self.pending.sort(key=lambda j: j.gpus)
            return
        if self.policy == "gandiva":
            # fifo over an explicit waiting list + time slicing: preempted jobs come back in `waiting` with
            # attempt > 0 and rejoin the list at the back; the list is served in order until the first job that
            # does not fit; then, while anything still waits, every job that has held its GPUs for a whole
            # quantum is asked to leave.
            for j in waiting:
                if j.job_id not in self.seen:
                    self.seen.add(j.job_id)
                    self.pending.append(j)
                elif j not in self.pending:
                    self.pending.append(j)
            for j in list(self.pending):
                if self._start(j, now):
                    self.pending.remove(j)
                else:
                    break
            if self.pending:
                for j in running:
                    if j.status == "RUNNING" and now - j.last_start >= self.quantum_s:
                        self._request_preempt(j, now)
            return
        # preemptive family: rank all runnable jobs, give GPUs by counting in that order
        runnable = waiting + running
        if self.policy == "sjf":
            runnable.sort(key=lambda j: (j.gpus, j.submit))
        elif self.policy == "shortest":
            runnable.sort(key=lambda j: ((j.iters - j.done) * self._iter_s(j), j.submit))
        elif self.policy == "dlas-gpu":
            # the simulator's dlas loop (sim/legacy/schedulers.py::dlas, reference run_sim.py:663-946) on explicit
            # queue lists: arrivals join the back of Q0; a running job whose attained GPU-time reached its
            # queue's limit moves to the back of the next queue (one level per pass); GPUs are handed out by
            # counting, queue by queue in list order; afterwards every queue keeps its running jobs in front
            # and its waiting ones, in order, behind them.
            if not self.queues:
                self.queues = [[] for _ in range(len(self.queue_limit) + 1)]
            for j in waiting:
                if j.job_id not in self.seen:
                    self.seen.add(j.job_id)
                    self.queues[0].append(j)
            live = {j.job_id for j in runnable}
            for q in self.queues:
                q[:] = [j for j in q if j.job_id in live]
            for qi in range(len(self.queues) - 1):
                for j in list(self.queues[qi]):
                    if j.status != "RUNNING" or j.q_id != qi:
                        continue
                    served = (j.attained + (now - j.last_start)) * j.gpus
                    if served >= self.queue_limit[qi]:
                        self.queues[qi].remove(j)
                        self.queues[qi + 1].append(j)
                        j.q_id = qi + 1
            runnable = [j for q in self.queues for j in q]
        elif self.policy == "gittins":
            # highest Gittins index of the attained GPU-time first (Tiresias 2D-Gittins; the reference's own
            # loop sorts ascending, run_sim.py:1082 — the simulator keeps that selectable, --gittins_desc)
            import torch
            from ..sim.legacy import gittins as gi
            if runnable:
                att = torch.tensor([int((j.attained + (now - j.last_start if j.status == "RUNNING" else 0.0))
                                        * j.gpus * 1000.0) for j in runnable], dtype=torch.int64)
                ranks = gi.gittins_ranks(self.gittins_table, att).tolist()
                for j, r in zip(runnable, ranks):
                    j.rank = float(r)
            runnable.sort(key=lambda j: (-j.rank, 0 if j.status != "WAITING" else 1, j.submit))
        else:
            raise NotImplementedError(self.policy)
        free = self.num_gpus
        chosen = []
        for j in runnable:
            if free >= j.gpus:
                chosen.append(j)
                free -= j.gpus
        chosen_ids = {j.job_id for j in chosen}
        for j in running:
            if j.job_id not in chosen_ids:
                self._request_preempt(j, now)
        if any(j.status == "PREEMPTING" for j in running):
            return                                 # start the winners once the losers have left their GPUs
        for j in chosen:
            if j.status == "WAITING":
                self._start(j, now)
        for q in self.queues:                       # dlas-gpu: waiting jobs go behind the running ones
            q[:] = [j for j in q if j.status != "WAITING"] + [j for j in q if j.status == "WAITING"]

    # -- main loop --------------------------------------------------------------------------------------------
    def run(self) -> List[ExecJob]:
        """Decisions are taken at the simulator's event instants only — a job arrives, or an attempt
        ends — plus, for dlas-gpu, on every poll (queue demotions happen between events) and, for gittins, every
        ``gittins_delta`` seconds (the simulator's extra wake-up)."""
        self.t0 = time.time()
        announced: set = set()
        while any(j.status not in ("DONE", "FAILED") for j in self.jobs):
            now = time.time() - self.t0
            changed = self._collect(now)
            arrivals = [j for j in self.jobs if j.submit <= now and j.job_id not in announced]
            announced.update(j.job_id for j in arrivals)
            tick = False
            if self.policy == "gittins" and now >= self._next_unit:      # the simulator's extra wake-up
                self._next_unit = (int(now / self.gittins_delta_s) + 1) * self.gittins_delta_s
                tick = True
            if changed or arrivals or tick or self.policy in ("dlas-gpu", "gandiva"):
                self._schedule(now)
            time.sleep(self.poll_s)
        return self.jobs

    def records(self) -> List[Dict]:
        out = []
        for j in self.jobs:
            iter_ms = [r["ranks"][0].get("iter_ms") for r in j.results if r["ranks"][0].get("iter_ms")]
            startup = [max(x.get("startup_s", 0.0) for x in r["ranks"]) for r in j.results]
            out.append({"job_id": j.job_id, "model": j.model, "num_gpu": j.gpus, "iterations": j.iters,
                        "submit": j.submit, "start": j.first_start, "end": j.end, "jct": j.jct(),
                        "status": j.status, "preempt": j.preempt, "attempts": len(j.results),
                        "iter_ms": sum(iter_ms) / len(iter_ms) if iter_ms else None,
                        "startup_s": sum(startup) / len(startup) if startup else None,
                        "snapshot_s": max([x.get("snapshot_s", 0.0) for r in j.results for x in r["ranks"]] or [0.0]),
                        "peak_mem_bytes": max([x.get("peak_mem_bytes", 0) for r in j.results for x in r["ranks"]] or [0]),
                        "gpu_util_avg": _mean([x.get("gpu_util_avg") for r in j.results for x in r["ranks"]]),
                        "gpu_util_max": max([x.get("gpu_util_max") or 0.0 for r in j.results for x in r["ranks"]] or [0.0]) or None})
        return out
