#!/usr/bin/env python
"""Warm-cache per-kernel device times of one training step via torch.profiler (CUPTI): complements the
ncu launch list, whose per-kernel cache flush inflates everything that normally hits the 126 MB L2.

    python bench/profile_step.py --model resnet50 --mode ps --batch 64
"""
import argparse
import collections
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--mode", default="ps")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--top", type=int, default=30)
    args = ap.parse_args()
    from tiresias_b200.models import zoo
    from tiresias_b200.parallel.trainer import Trainer
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    spec = zoo.resolve(args.model)
    tr = Trainer(spec, batch_size=args.batch, mode=args.mode, use_graph=False, device=dev)
    x, y = zoo.synthetic_batch(spec, args.batch)
    for _ in range(5):
        tr.step(x, y)
    tr.loss()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(args.steps):
            tr.step(x, y)
        tr.loss()
        torch.cuda.synchronize()
    agg = collections.defaultdict(lambda: [0, 0.0])
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA and ev.device_time > 0:
            agg[ev.name[:90]][0] += 1
            agg[ev.name[:90]][1] += ev.device_time
    tot = sum(v[1] for v in agg.values())
    print(f"# {args.model} mode={args.mode} batch={args.batch}: {tot / args.steps / 1e3:.3f} ms of kernel time per step, "
          f"{sum(v[0] for v in agg.values()) // args.steps} launches per step (warm cache, CUPTI)")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:args.top]:
        print(f"{t / args.steps / 1e3:8.3f} ms {100 * t / tot:5.1f}% n={n // args.steps:4d} {k}")
    tr.close()


if __name__ == "__main__":
    main()
