#!/usr/bin/env python
"""Synthetic README-schema trace sized for a real replay on one 8xB200 box: mixed models and GPU counts,
Poisson arrivals, iteration counts that give jobs of a few seconds each.

    python scripts/make_exec_trace.py out.csv --jobs 64 --seed 0 --mean-gap 1.5
"""
import argparse
import csv

import numpy as np

MODELS = [("resnet50", 0.4), ("vgg16", 0.2), ("transformer", 0.2), ("alexnet", 0.1), ("resnet101", 0.1)]
GPUS = [(1, 0.45), (2, 0.25), (4, 0.2), (8, 0.1)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--jobs", type=int, default=64)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--mean-gap", type=float, default=1.5, help="mean inter-arrival in seconds")
    ap.add_argument("--min-iters", type=int, default=60)
    ap.add_argument("--max-iters", type=int, default=400)
    a = ap.parse_args()
    rng = np.random.default_rng(a.seed)
    t = np.cumsum(rng.exponential(a.mean_gap, size=a.jobs))
    with open(a.out, "w", newline="") as fh:
        w = csv.writer(fh, lineterminator="\n")
        w.writerow(["job_id", "num_gpu", "submit_time", "iterations", "model_name", "duration", "interval"])
        for i in range(a.jobs):
            model = rng.choice([m for m, _ in MODELS], p=[p for _, p in MODELS])
            gpus = int(rng.choice([g for g, _ in GPUS], p=[p for _, p in GPUS]))
            iters = int(np.exp(rng.uniform(np.log(a.min_iters), np.log(a.max_iters))))
            w.writerow([i, gpus, round(float(t[i]), 3), iters, model, max(1, iters // 20), 0])


if __name__ == "__main__":
    main()
