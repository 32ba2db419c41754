"""Logs of the event-driven schedulers (reference: log.py:10-352, D13).

``job.csv`` (and its README-named twin ``jobs.csv``) gets one row per completed job:
``[event_time, job_id, num_gpu, submit_time, start_time, end_time, end-start, JCT = end-submit,
duration, pending_time, preempt, (resume,) promote]`` — ``resume`` only under ``--scheme count``.
``cluster.csv`` gets one row per event:
``[time, idle_node, busy_node, full_node, idle_gpu, busy_gpu, pending_job, running_job,
completed_job]`` (the gandiva variant appends the node-set counts). With a real placement scheme the
per-node files are filled as the reference's commented-out accounting intended (log.py:171-195):
``cpu.csv`` busy CPUs, ``gpu.csv`` busy GPUs and ``network.csv`` in/out MB per node, ``memory.csv``
max / 99th / 95th / median used memory over the nodes that use any.
"""
from __future__ import annotations

import csv
import math
import os
from typing import List

CLUSTER_HEADER = ["time", "idle_node", "busy_node", "full_node", "idle_gpu", "busy_gpu", "pending_job",
                  "running_job", "completed_job"]
GANDIVA_EXTRA = ["len_g1", "len_g2", "len_g4", "len_g8", "len_g16", "len_g32", "len_g64"]
JOB_HEADER = ["time", "job_id", "num_gpu", "submit_time", "start_time", "end_time", "executed_time", "JCT",
              "duration", "pending_time", "preempt", "promote"]
JOB_HEADER_COUNT = JOB_HEADER[:-1] + ["resume", "promote"]
GPU_DEMANDS = [1, 2, 4, 8, 12, 16, 24, 32]


class TLog:
    def __init__(self, path: str, flags, cluster):
        self.log_path = path
        self.flags = flags
        self.cluster = cluster
        self.is_count = flags.scheme in ("count", "buddy")
        self.finished_rows: List[list] = []
        self._files = {}

    def _open(self, name: str, header: List[str]):
        fh = open(os.path.join(self.log_path, name), "w+", newline="")
        w = csv.writer(fh)
        w.writerow(header)
        self._files[name] = (fh, w)

    def init_log(self) -> None:
        c = self.cluster
        header = list(CLUSTER_HEADER)
        if self.flags.scheme == "gandiva" or self.flags.schedule == "gandiva-legacy":
            header = header[:4] + ["fra_gpu"] + header[5:] + GANDIVA_EXTRA
        self._open("cluster.csv", header)
        if not self.is_count:
            self._open("cpu.csv", ["time"] + ["cpu" + str(i) for i in range(c.num_node)])
            self._open("gpu.csv", ["time"] + ["gpu" + str(i) for i in range(c.num_node)])
            titles = ["time"]
            for i in range(c.num_node):
                titles += ["in" + str(i), "out" + str(i)]
            self._open("network.csv", titles)
            self._open("memory.csv", ["time", "max", "99th", "95th", "med"])
        if self.flags.schedule == "gpu-demands":
            job_header = ["time"] + ["%d-GPU" % g for g in GPU_DEMANDS]
        else:
            job_header = JOB_HEADER_COUNT if self.is_count else JOB_HEADER
        self._open("job.csv", job_header)
        self._open("jobs.csv", job_header)

    def _row(self, name: str, row: list) -> None:
        fh, w = self._files[name]
        w.writerow(row)
        fh.flush()

    def close(self) -> None:
        for fh, _ in self._files.values():
            fh.close()
        self._files.clear()

    # -- cluster rows -------------------------------------------------------------------------------------
    def _job_counts(self, jobs):
        running = pending = completed = 0
        for j in jobs.job_list:
            if j.status == "RUNNING":
                running += 1
            elif j.status == "PENDING":
                pending += 1
            elif j.status == "END":
                completed += 1
        return pending, running, completed

    def checkpoint(self, jobs, event_time) -> None:
        c = self.cluster
        idle_node = busy_node = full_node = idle_gpu = busy_gpu = 0
        if not self.is_count:
            cpu, gpu, net, mem = [event_time], [event_time], [event_time], []
            for node in c.nodes():
                free = node.free_gpus
                idle_gpu += free
                busy_gpu += node.num_gpu - free
                if free == node.num_gpu:
                    idle_node += 1
                elif free > 0:
                    busy_node += 1
                else:
                    full_node += 1
                cpu.append(node.num_cpu - node.free_cpus)
                gpu.append(node.num_gpu - free)
                net += [node.network_in, node.network_out]
                used_mem = self.flags.mem_p_node - node.free_mem + 2
                if used_mem > 2:
                    mem.append(used_mem)
            self._row("cpu.csv", cpu)
            self._row("gpu.csv", gpu)
            self._row("network.csv", net)
            if mem:
                n = len(mem)
                if n == 1:
                    i99 = i95 = imed = 0
                else:
                    i99 = min(int(math.ceil(n * 0.01)), n - 1)
                    i95 = min(int(math.ceil(n * 0.05)), n - 1)
                    imed = (n - 1) // 2
                top = sorted(mem, reverse=True)
                self._row("memory.csv", [event_time, round(top[0], 1), round(top[i99], 1),
                                         round(top[i95], 1), round(top[imed], 1)])
        else:
            if self.flags.schedule == "dlas-gpu-pack":
                idle_gpu = sum(1 for g in c.gpu_list if g == 1)
                busy_gpu = len(c.gpu_list) - idle_gpu
            else:
                idle_gpu = c.free_gpu
                busy_gpu = c.num_gpu - c.free_gpu
            busy_node = int(math.ceil(busy_gpu / c.num_gpu_p_node))
            full_node = busy_node
            idle_node = int(c.num_node - busy_node)
        pending, running, completed = self._job_counts(jobs)
        self._row("cluster.csv", [event_time, idle_node, busy_node, full_node, idle_gpu, busy_gpu,
                                  pending, running, completed])

    def checkpoint_multi_dlas_gpu(self, jobs, event_time) -> None:
        c = self.cluster
        idle_gpu = sum(g.free_gpu for g in jobs.gpu_job.values())
        busy_gpu = c.num_gpu - idle_gpu
        busy_node = int(math.ceil(busy_gpu / c.num_gpu_p_node))
        pending, running, completed = self._job_counts(jobs)
        self._row("cluster.csv", [event_time, int(c.num_node - busy_node), busy_node, busy_node,
                                  int(idle_gpu), int(busy_gpu), pending, running, completed])

    def gandiva_checkpoint(self, jobs, event_time, idle_node, busy_gpu, frag_gpu, pending_job, running_job,
                           len_g1, len_g2, len_g4, len_g8, len_g16, len_g32, len_g64) -> None:
        busy_node = self.cluster.num_node - idle_node
        self._row("cluster.csv", [event_time, idle_node, busy_node, 0, frag_gpu, busy_gpu, pending_job,
                                  running_job, len(jobs.completed_jobs), len_g1, len_g2, len_g4, len_g8,
                                  len_g16, len_g32, len_g64])

    def checkpoint_gpu_demands(self, jobs, event_time) -> None:
        row = [event_time] + [g * jobs.gpu_job.get(g, 0) for g in GPU_DEMANDS]
        self._row("job.csv", row)
        self._row("jobs.csv", row)

    # -- job rows -------------------------------------------------------------------------------------------
    def job_complete(self, job, event_time) -> None:
        job.end_time = event_time
        executed = job.end_time - job.start_time
        jct = int(job.end_time - job.submit_time)
        row = [event_time, job.job_id, job.num_gpu, job.submit_time, job.start_time, job.end_time, executed,
               jct, job.duration, job.pending_time, job.preempt]
        row += [job.resume, job.promote] if self.is_count else [job.promote]
        self.finished_rows.append(row)
        self._row("job.csv", row)
        self._row("jobs.csv", row)
