"""Worker pool: one long-lived process per GPU that trains whatever job the controller assigns to it.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 \
        -m tiresias_b200.exec.worker --ctl-port P [--loopback]

Start-up: NCCL (or gloo with ``--loopback``) world of G ranks, then one *gloo* group per aligned
power-of-two GPU block (:func:`tiresias_b200.exec.buddy.all_blocks`) — the bootstrap channel over which a
job's ranks exchange their CUDA-IPC handles (``SymmetricWorld``) and agree on preemption. With
``--with-nccl`` the same blocks also get NCCL groups for the ``nccl`` baseline mode.

Protocol (a ``TCPStore`` served by the controller):
  controller -> rank r : key ``cmd/<r>/<seq>`` = JSON {op: run|exit, job, attempt, model, batch, iters,
                         done, mode, block: [start, size], resume: bool, ckpt_dir, check_every}
  controller -> job    : key ``preempt/<job>/<attempt>`` (existence = please stop)
  rank -> controller   : key ``res/<job>/<attempt>/<rank_in_block>`` = JSON result

Packing (reference: up to 3 tasks per device, infra/device.py:49-59): a ``run`` command with ``"shared": true``
(single-GPU jobs only) is executed in a thread of its own inside the GPU's worker process — same CUDA context,
its own streams — while the command loop keeps listening, so up to three jobs really run concurrently on one
B200 (kernels of different streams overlap; no context time-slicing). Exclusive jobs run inline as before.

A run trains ``iters - done`` iterations of the job's model on the block's GPUs with the requested
communication mode (``ps`` fused push/pull kernels, ``ring`` in-kernel all-reduce, ``nccl`` baseline).
Every ``check_every`` iterations the ranks agree (max over a gloo all-reduce) whether a preemption was
requested; if so each rank snapshots its PS shard (fp32 master + momentum) to ``ckpt_dir`` — host memory
when that is /dev/shm — and reports how far it got. A resumed run loads those shards, on whatever block
it is given next.
"""
from __future__ import annotations

import argparse
import json
import os
import time
import traceback
from typing import Dict, Tuple

import torch
import torch.distributed as dist

from .buddy import all_blocks


TENANTS = [0]        # single-GPU jobs currently sharing this worker's GPU (threads of this process)


def _make_trainer(cmd: Dict, rank_in_block: int, size: int, group, nccl_group, loopback: bool, device):
    if loopback:
        from .loopback import LoopbackTrainer
        tr = LoopbackTrainer(cmd["model"], rank_in_block, size, group, step_sleep=cmd.get("step_sleep", 0.0))
        table = cmd.get("interference")
        if table and cmd.get("shared"):
            # emulate co-location on CPU: a step takes slowdown[tenants - 1] times its solo time
            tr.slowdown = lambda: float(table[min(max(TENANTS[0], 1), len(table)) - 1])
        return tr
    from ..models import zoo
    from ..parallel.symm import SymmetricWorld
    from ..parallel.trainer import Trainer
    spec = zoo.resolve(cmd["model"])
    world = SymmetricWorld(rank_in_block, size, device, group=group)
    return Trainer(spec, cmd["batch"], mode=cmd.get("mode", "ps"), world=world, device=device,
                   use_graph=bool(cmd.get("graph", True)), comm_group=nccl_group,
                   bucket_mb=float(cmd.get("bucket_mb", 8.0)))


def run_job(cmd: Dict, rank: int, store, groups, nccl_groups, loopback: bool, device) -> Dict:
    start, size = cmd["block"]
    r = rank - start
    grp = groups[(start, size)]
    t_begin = time.time()
    shared = bool(cmd.get("shared")) and size == 1
    if shared:
        TENANTS[0] += 1
    try:
        return _run_job_body(cmd, rank, store, groups, nccl_groups, loopback, device, r, size, grp, t_begin)
    finally:
        if shared:
            TENANTS[0] -= 1


def _run_job_body(cmd, rank, store, groups, nccl_groups, loopback, device, r, size, grp, t_begin) -> Dict:
    start = cmd["block"][0]
    trainer = _make_trainer(cmd, r, size, grp, nccl_groups.get((start, size)), loopback, device)
    ckpt = os.path.join(cmd["ckpt_dir"], f"{cmd['job']}_{r}.pt")
    if cmd.get("resume"):
        trainer.load_state_dict(torch.load(ckpt))
    if not loopback:
        from ..models import zoo
        x, y = zoo.synthetic_batch(zoo.resolve(cmd["model"]), cmd["batch"], seed=rank)
        trainer.step(x, y)                     # loads the batch; the remaining steps reuse it on the device
        trainer.loss()
        done0 = 1
    else:
        done0 = 0
    done, iters = int(cmd.get("done", 0)) + done0, int(cmd["iters"])
    check_every = max(1, int(cmd.get("check_every", 10)))
    preempt_key = f"preempt/{cmd['job']}/{cmd['attempt']}"
    first = done
    preempted = False
    flag = torch.zeros(1, dtype=torch.int32)
    # The first chunk absorbs the eager warm-up steps and the CUDA-graph capture; the steady-state
    # iteration time is measured from the end of that chunk on, and everything else this run cost
    # (model build, IPC exchange, warm-up, capture) is reported as one fixed overhead, so that
    #     run time = overhead_s + iterations * iter_ms        holds for jobs of any length.
    warm_chunk = max(check_every, 8)
    t_steady = None
    steady_from = done
    util_sampler = None
    while done < iters:
        upto = min(iters, done + (warm_chunk if t_steady is None else check_every))
        for _ in range(upto - done):
            trainer.step_resident()
        done = upto
        if not loopback and trainer.check_status() != 0:
            # an in-kernel wait ran out of its spin budget (a peer is gone): stop now instead of training on
            # half-updated shards; the controller marks the job failed and frees its GPUs
            raise RuntimeError(f"in-kernel wait timed out (status {trainer.check_status()}) at iteration {done}")
        if t_steady is None:
            trainer.synchronize()
            t_steady, steady_from = time.time(), done
            if not loopback and torch.cuda.is_available():          # K7: measured utilisation of this job
                from ..utils.clocks import UtilSampler
                util_sampler = UtilSampler(device.index or 0).start()
        if done < iters:
            flag[0] = 1 if (r == 0 and store.check([preempt_key])) else 0
            if size > 1:
                dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=grp)
            if int(flag[0]):
                preempted = True
                break
    trainer.synchronize()
    t_done = time.time()
    util = util_sampler.stop() if util_sampler is not None else {"gpu_util_avg": None, "gpu_util_max": None}
    run_s = t_done - t_begin
    steady_iters = done - steady_from
    if t_steady is not None and steady_iters >= 5:
        iter_ms = (t_done - t_steady) / steady_iters * 1e3
    else:
        iter_ms = None
    iters_run = done - first + done0
    startup_s = run_s - iters_run * iter_ms / 1e3 if iter_ms is not None else run_s
    status = trainer.check_status()
    snapshot_s = 0.0
    if preempted:
        t0 = time.time()
        os.makedirs(cmd["ckpt_dir"], exist_ok=True)
        torch.save(trainer.state_dict(), ckpt)
        snapshot_s = time.time() - t0
    peak = torch.cuda.max_memory_allocated(device) if (not loopback and torch.cuda.is_available()) else 0
    loss = trainer.loss()
    trainer.close()
    return {"job": cmd["job"], "attempt": cmd["attempt"], "rank": r, "status": "preempted" if preempted else "done",
            "done": done, "iters_run": iters_run, "run_s": run_s, "startup_s": startup_s,
            "iter_ms": iter_ms, "snapshot_s": snapshot_s, "kernel_status": status,
            "peak_mem_bytes": int(peak), "gpu_util_avg": util["gpu_util_avg"], "gpu_util_max": util["gpu_util_max"],
            "loss": loss, "t_end": time.time()}


def _warm_up(device) -> None:
    """Pay the one-off costs (CUDA context, cuDNN / cuBLAS handles, allocator pools) before the first job,
    so that every job's start-up time is the same few tens of milliseconds."""
    x = torch.randn(8, 16, 32, 32, device=device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    conv = torch.nn.Conv2d(16, 16, 3, padding=1).to(device).to(torch.bfloat16).to(memory_format=torch.channels_last)
    y = conv(x)
    y.float().sum().backward()
    a = torch.randn(256, 256, device=device, dtype=torch.bfloat16)
    (a @ a).sum().item()
    torch.nn.functional.scaled_dot_product_attention(a.view(1, 4, 256, 64), a.view(1, 4, 256, 64), a.view(1, 4, 256, 64))
    torch.cuda.synchronize(device)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--ctl-port", type=int, required=True)
    ap.add_argument("--ctl-host", default="127.0.0.1")
    ap.add_argument("--loopback", action="store_true", help="CPU tensors over gloo (no GPU needed)")
    ap.add_argument("--with-nccl", action="store_true", help="also build NCCL groups per block (mode nccl)")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    device = None
    if args.loopback:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    else:
        torch.cuda.set_device(local)
        device = torch.device("cuda", local)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
        from ..ops import native
        native.lib()
        _warm_up(device)
    groups: Dict[Tuple[int, int], object] = {}
    nccl_groups: Dict[Tuple[int, int], object] = {}
    for blk in all_blocks(world):
        ranks = list(range(blk[0], blk[0] + blk[1]))
        groups[blk] = dist.new_group(ranks, backend="gloo")
        if args.with_nccl and not args.loopback and blk[1] > 1:
            nccl_groups[blk] = dist.new_group(ranks, backend="nccl")
    store = dist.TCPStore(args.ctl_host, args.ctl_port, is_master=False, timeout=__import__("datetime").timedelta(days=1))
    store.set(f"ready/{rank}", "1")
    import threading

    def run_and_report(cmd: Dict, own_client: bool = False) -> None:
        # a tenant thread talks to the controller over its own connection: the command loop's client is parked
        # in a blocking get() most of the time
        st = store
        if own_client:
            st = dist.TCPStore(args.ctl_host, args.ctl_port, is_master=False,
                               timeout=__import__("datetime").timedelta(days=1))
        try:
            if device is not None:
                torch.cuda.set_device(device)                # per thread
            res = run_job(cmd, rank, st, groups, nccl_groups, args.loopback, device)
        except Exception as exc:                             # report instead of dying: the pool survives
            res = {"job": cmd.get("job"), "attempt": cmd.get("attempt"), "rank": rank - cmd["block"][0],
                   "status": "failed", "error": f"{type(exc).__name__}: {exc}", "trace": traceback.format_exc()[-1500:],
                   "t_end": time.time()}
        st.set(f"res/{cmd['job']}/{cmd['attempt']}/{res['rank']}", json.dumps(res))

    seq = 0
    tenants = []
    while True:
        raw = store.get(f"cmd/{rank}/{seq}")              # blocks until the controller posts it
        seq += 1
        cmd = json.loads(raw)
        if cmd["op"] == "exit":
            break
        tenants = [t for t in tenants if t.is_alive()]
        if cmd.get("shared") and cmd["block"][1] == 1:
            # a packed single-GPU job: runs beside this GPU's other tenants, the loop keeps listening
            t = threading.Thread(target=run_and_report, args=(cmd, True), daemon=True)
            t.start()
            tenants.append(t)
        else:
            run_and_report(cmd)
    for t in tenants:
        t.join()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
