#!/usr/bin/env python
"""Headline benchmark: ResNet-50 parameter-server/worker synchronous SGD, samples/s over N B200s.

BASELINE.json config 3: "ResNet-50 PS/worker sync SGD executed on 1/2/4/8 B200: fused push/pull
vs NCCL baseline". Weak scaling (per-GPU batch fixed), bf16 compute with fp32 master weights on
the PS shards, synthetic uint8 images and random-init weights (no network for datasets).

  python bench.py --gpus 1 --steps 20 --warmup 5
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port P bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0. ``value`` is device-timed (CUDA events, max over ranks) with the
batch resident on the GPU; ``e2e`` goes through the public ``Trainer.step`` API including the
pinned host->device copy of every step's inputs and the device->host read of every step's loss.
Two NCCL arms are measured in the same process right after:
``nccl_baseline``      the stock PyTorch + NCCL PS/worker job BASELINE.json prescribes (nn.BatchNorm2d, cuDNN /
                       cuBLAS, eager, per-parameter gradient accumulation, torch optimizer ops);
``nccl_same_compute``  the SAME fused network, gradient packing, optimizer kernel and CUDA graph as ours —
                       only the communication differs (NCCL reduce_scatter + all_gather instead of the fused
                       push/pull kernel). ours / nccl_same_compute is what the in-kernel push/pull buys;
                       nccl_same_compute / nccl_baseline is what the compute kernels and the graph buy.
``sim`` carries the other half of BASELINE's metric (makespan per scheduler): the UNMODIFIED reference
simulator from ``baseline/_ref`` driven through its own ``run_sim.py`` next to this simulator on the same
trace and seed — wall time of both, makespans, and whether the two cluster.csv files are byte-identical.
``--impl reference`` runs only the reference: it has no execution path, so it reports the samples/s metric as
unavailable and emits the same ``sim`` keys.
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import sys
import time


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def parse_args() -> argparse.Namespace:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "nccl", "ring"])
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batch", type=int, default=256, help="per-GPU batch (weak scaling)")
    ap.add_argument("--bucket-mb", type=float, default=8.0)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--skip-nccl", action="store_true", help="do not run the NCCL baseline arms")
    ap.add_argument("--skip-sim", action="store_true", help="do not run the simulator comparison")
    ap.add_argument("--extra-models", default="vgg16:128,transformer_big:32",
                    help="communication-bound jobs reported next to the headline as model:per-GPU-batch pairs "
                         "(VGG-16 528 MB and transformer_big 1.1 GB of gradients); '' = none")
    return ap.parse_args()


def sim_block(ours: bool, reference: bool):
    """Simulator comparison (rank 0 only, CPU): see tiresias_b200/utils/refsim.py."""
    try:
        from tiresias_b200.utils import refsim
        return refsim.compare(ours=ours, reference=reference)
    except Exception as e:  # noqa: BLE001 - context for the headline, never a reason to lose it
        return {"error": f"{type(e).__name__}: {e}"[:300]}


def reference_arm(args: argparse.Namespace) -> None:
    # Install outcome: `pip install --target baseline/_ref /root/reference` fails with
    # "Neither 'setup.py' nor 'pyproject.toml' found", so baseline/_ref is an untouched copy of the tree. The
    # reference has no code path that executes a model (CPU-only scheduling simulator), so the samples/s
    # metric cannot be produced by it; what it CAN produce — makespan per scheduler through its own
    # run_sim.py — is measured here.
    line = {
        "impl": "reference",
        "unavailable": "reference is a CPU-only scheduling simulator with no training/execution path: it cannot "
                       "produce samples/s (and has no setup.py/pyproject.toml; baseline/_ref is an unmodified copy)",
        "metric": f"{args.model}_ps_sync_sgd_samples_per_s", "value": None, "n_gpus": args.gpus,
    }
    if int(os.environ.get("RANK", "0")) == 0:
        if not args.skip_sim:
            line["sim"] = sim_block(ours=False, reference=True)
        print(json.dumps(line))


def init_dist(args: argparse.Namespace):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun for N>1")
    torch.cuda.set_device(local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(_free_port()))
    dist.init_process_group("nccl", rank=rank, world_size=world,
                            device_id=torch.device("cuda", local_rank))
    return rank, local_rank, world


def timed_region(fn_step, steps: int, device, sampler=None):
    """EXACTLY `steps` calls of fn_step between barrier+synchronize brackets; returns the
    device-side elapsed milliseconds (max over ranks) and the wall time."""
    import torch
    import torch.distributed as dist
    dist.barrier()
    torch.cuda.synchronize(device)
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if sampler is not None:
        sampler.start()
    t0 = time.perf_counter()
    start.record(torch.cuda.current_stream(device))
    fn_step(steps)
    end.record(torch.cuda.current_stream(device))
    torch.cuda.synchronize(device)
    dist.barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop() if sampler is not None else None
    ms = torch.tensor([start.elapsed_time(end)], device=device)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()), wall, clocks


def run_arm(mode: str, args: argparse.Namespace, rank: int, local_rank: int, world: int, same_compute: bool = False,
            model: str = None, batch: int = None, steps: int = None, phases: bool = False):
    import torch
    from tiresias_b200.models import zoo
    from tiresias_b200.ops import native
    from tiresias_b200.parallel.trainer import Trainer
    from tiresias_b200.utils.clocks import ClockSampler

    device = torch.device("cuda", local_rank)
    model = model or args.model
    batch = batch or args.batch
    steps = steps or args.steps
    spec = zoo.resolve(model)
    trainer = Trainer(spec, batch, mode=mode, bucket_mb=args.bucket_mb,
                      use_graph=not args.no_graph, device=device, nccl_same_compute=same_compute)
    # distinct pinned host batches so every step's H2D copy moves fresh bytes
    host = [zoo.synthetic_batch(spec, batch, pinned=True, seed=1000 * rank + i) for i in range(4)]

    cur = torch.cuda.current_stream(device)

    def resident(k: int) -> None:
        cur.wait_stream(trainer.stream)
        for _ in range(k):
            trainer.step_resident()
        cur.wait_stream(trainer.stream)
        cur.wait_stream(trainer.comm_stream)

    losses = []

    def e2e(k: int) -> None:
        # the public training loop: every step's batch crosses PCIe from pinned memory (the copy of batch
        # i+1 overlaps step i, Trainer.fit) and every step's loss is read back to the host
        cur.wait_stream(trainer.stream)
        for loss in trainer.fit(host[i % len(host)] for i in range(k)):
            losses.append(loss)
        cur.wait_stream(trainer.stream)
        cur.wait_stream(trainer.comm_stream)
        cur.wait_stream(trainer.copy_stream)

    # warm-up through the public API: covers allocator growth, cuDNN autotune and graph capture
    e2e(max(args.warmup, 5 if trainer.use_graph and (mode != "nccl" or same_compute) else args.warmup))
    if trainer.check_status() != 0:
        raise SystemExit("in-kernel wait timed out during warm-up")

    launches0 = trainer.native_launches
    ms, wall, clocks = timed_region(resident, steps, device, ClockSampler(local_rank))
    launches = trainer.native_launches - launches0
    ms_e2e, wall_e2e, _ = timed_region(e2e, steps, device)
    status = trainer.check_status()
    phase_ms = None
    if phases:
        # eager per-phase device times (max over ranks): how much communication is left exposed after backward
        try:
            phase_ms = {k: round(v, 3) for k, v in trainer.profile_phases(*host[0], steps=3).items()}
        except Exception as e:  # noqa: BLE001
            phase_ms = {"error": f"{type(e).__name__}: {e}"[:200]}
    out = {
        "ms_per_step": ms / steps,
        "value": batch * world * steps / (ms / 1e3),
        "e2e_ms_per_step": ms_e2e / steps,
        "e2e_value": batch * world * steps / (ms_e2e / 1e3),
        "phases_ms": phase_ms, "nvls": bool(getattr(trainer, "nvls", False)),
        "grad_mb": round(sum(p.numel() for p in trainer.model.parameters()) * 2 / 1e6, 1),
        "pull_layers": sum(1 for b in trainer.buckets if b.pull_layer is not None),
        "wall_s": wall, "clocks": clocks, "gpu_launches": int(launches),
        "h2d_bytes_per_step": trainer.h2d_bytes, "d2h_bytes_per_step": trainer.d2h_bytes,
        "loss_last": losses[-1] if losses else None, "status": status,
        "buckets": len(trainer.buckets), "graph": bool(trainer._graph is not None),
        "params": sum(p.numel() for p in trainer.model.parameters()),
    }
    trainer.close()
    del trainer
    torch.cuda.empty_cache()
    return out


def main() -> None:
    args = parse_args()
    if args.impl == "reference":
        reference_arm(args)
        return
    import torch
    import torch.distributed as dist
    rank, local_rank, world = init_dist(args)
    mode = {"ours": "ps", "nccl": "nccl", "ring": "ring"}[args.impl]
    sim = None
    if rank == 0 and args.impl == "ours" and not args.skip_sim:
        sim = sim_block(ours=True, reference=True)     # CPU only; the other ranks wait at the first barrier
    res = run_arm(mode, args, rank, local_rank, world, phases=True)
    base, fair, base_error = None, None, None
    if args.impl == "ours" and not args.skip_nccl:
        # the PyTorch + NCCL arms are context for the headline number, never a reason to lose it
        try:
            fair = run_arm("nccl", args, rank, local_rank, world, same_compute=True)
        except Exception as e:  # noqa: BLE001
            base_error = f"nccl_same_compute: {type(e).__name__}: {e}"[:300]
        try:
            base = run_arm("nccl", args, rank, local_rank, world)
        except Exception as e:  # noqa: BLE001
            base_error = f"{type(e).__name__}: {e}"[:300]
    extra = []
    if args.impl == "ours" and not args.skip_nccl and args.extra_models:
        # where communication is visible: same two arms (ours vs same-compute NCCL) on the big-gradient jobs
        for item in args.extra_models.split(","):
            name, _, b = item.partition(":")
            row = {"model": name, "per_gpu_batch": int(b or args.batch)}
            try:
                k = max(5, args.steps // 2)
                o = run_arm("ps", args, rank, local_rank, world, model=name, batch=row["per_gpu_batch"], steps=k, phases=True)
                f = run_arm("nccl", args, rank, local_rank, world, same_compute=True, model=name,
                            batch=row["per_gpu_batch"], steps=k)
                row.update({"gradient_mb_bf16": o["grad_mb"], "ours_ms_per_step": round(o["ms_per_step"], 3),
                            "ours_value": round(o["value"], 1), "nccl_same_compute_ms_per_step": round(f["ms_per_step"], 3),
                            "speedup_device": round(f["ms_per_step"] / o["ms_per_step"], 4),
                            "speedup_e2e": round(o["e2e_value"] / f["e2e_value"], 4), "phases_ms": o["phases_ms"],
                            "nvls": o["nvls"], "pull_layers": o["pull_layers"], "steps": k})
            except Exception as e:  # noqa: BLE001
                row["error"] = f"{type(e).__name__}: {e}"[:300]
            extra.append(row)
    if rank == 0:
        line = {
            "metric": f"{args.model}_ps_sync_sgd_samples_per_s",
            "value": round(res["value"], 2), "unit": "samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(res["ms_per_step"], 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic uint8 images, random-init weights",
            "impl": {"ps": "ours", "nccl": "nccl-baseline", "ring": "ours-ring"}[mode],
            "config": {
                "model": args.model, "global_batch": args.batch * world, "per_gpu_batch": args.batch,
                "seq_len": None, "parallelism": f"ps-worker dp{world} (1 worker + 1 PS shard per GPU)",
                "optimizer": "sgd momentum 0.9 wd 1e-4, fp32 master on PS shards",
                "buckets": res["buckets"], "cuda_graph": res["graph"],
                "l2": "per-step working set (activations + weights + grads, GBs) >> 126 MB L2; "
                      "e2e copies a fresh pinned batch every step (prefetched one step ahead); no explicit flush",
                "timing": "CUDA events on the launching stream, max over ranks",
            },
            "clocks": res["clocks"],
            "e2e": {"value": round(res["e2e_value"], 2), "unit": "samples/s",
                    "ms_per_step": round(res["e2e_ms_per_step"], 4),
                    "h2d_bytes_per_step": res["h2d_bytes_per_step"],
                    "d2h_bytes_per_step": res["d2h_bytes_per_step"]},
            "gpu_launches": res["gpu_launches"],
            "loss_last": res["loss_last"], "status": res["status"],
            "phases_ms": res["phases_ms"], "nvls": res["nvls"], "pull_layers": res["pull_layers"],
        }
        if extra:
            line["communication_bound_jobs"] = extra
        if base is not None:
            line["nccl_baseline"] = {
                "what": "stock PyTorch network (nn.BatchNorm2d, cuDNN/cuBLAS), eager, NCCL reduce_scatter + torch SGD + all_gather",
                "value": round(base["value"], 2), "ms_per_step": round(base["ms_per_step"], 4),
                "e2e_value": round(base["e2e_value"], 2),
                "speedup_device": round(res["value"] / base["value"], 4),
                "speedup_e2e": round(res["e2e_value"] / base["e2e_value"], 4),
            }
        if fair is not None:
            line["nccl_same_compute"] = {
                "what": "same fused network, gradient packing, SGD kernel and CUDA graph as ours; only the collectives are NCCL",
                "value": round(fair["value"], 2), "ms_per_step": round(fair["ms_per_step"], 4),
                "e2e_value": round(fair["e2e_value"], 2), "cuda_graph": fair["graph"],
                "speedup_device": round(res["value"] / fair["value"], 4),
                "speedup_e2e": round(res["e2e_value"] / fair["e2e_value"], 4),
            }
        if base_error is not None:
            line["nccl_error"] = base_error
        if sim is not None:
            line["sim"] = sim
        print(json.dumps(line))
        sys.stdout.flush()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
