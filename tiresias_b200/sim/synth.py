"""Synthetic traces (the reference ships none; ``data/*.csv`` is git-ignored there).

``live_trace`` follows the recipe of SURVEY.md §4 item 2; ``tiresias_trace`` writes the README schema
with model names drawn from the model table.
"""
from __future__ import annotations

import csv
from typing import Optional, Sequence

import numpy as np

from .models import m_names
from .trace import LIVE_COLUMNS, README_COLUMNS

GPU_CHOICES = (1, 1, 1, 2, 2, 4, 8, 16)


def live_trace(path: str, num_jobs: int, seed: int = 0, gap: float = 40000.0,
               oversized: Optional[int] = None, interactive_every: int = 0) -> str:
    """Write a live-schema csv. ``oversized`` appends one job with that many GPUs (to pin the
    "abandoned at the end" behaviour); ``interactive_every`` sprinkles rows of type interactive."""
    rng = np.random.default_rng(seed)
    t = np.cumsum(rng.uniform(0, gap, size=num_jobs)) + 1.5e9
    rows = []
    for i in range(num_jobs):
        gpus = int(rng.choice(GPU_CHOICES))
        per = 1 if gpus <= 8 else 8
        avg = float(rng.uniform(5, 70))
        mx = min(100.0, avg + float(rng.uniform(5, 30)))
        kind = "interactive" if interactive_every and i % interactive_every == interactive_every - 1 \
            else "noninteractive"
        rows.append([kind, float(t[i]), int(rng.integers(20, 400)), per, gpus, avg, mx,
                     int(rng.uniform(1, 12) * (1 << 30)), int(rng.uniform(0.5, 1) * (1 << 30))])
    if oversized:
        rows.append(["noninteractive", float(t[-1] + 1), 30, 8, oversized, 50.0, 80.0, 4 << 30, 1 << 30])
    with open(path, "w", newline="") as fh:
        w = csv.writer(fh, lineterminator="\n")
        w.writerow(LIVE_COLUMNS)
        w.writerows(rows)
    return path


def tiresias_trace(path: str, num_jobs: int, seed: int = 0, mean_gap: float = 120.0,
                   models: Sequence[str] = tuple(m_names), max_duration: int = 20000) -> str:
    rng = np.random.default_rng(seed)
    submit = np.cumsum(rng.exponential(mean_gap, size=num_jobs)).astype(int)
    deli = " " if path.endswith(".txt") else ","
    with open(path, "w", newline="") as fh:
        w = csv.writer(fh, delimiter=deli, lineterminator="\n")
        w.writerow(README_COLUMNS)
        for i in range(num_jobs):
            gpus = int(rng.choice(GPU_CHOICES))
            dur = int(min(max_duration, max(60, rng.lognormal(6.5, 1.3))))
            w.writerow([i, gpus, int(submit[i]), int(dur * rng.uniform(2, 6)), str(rng.choice(models)),
                        dur, int(rng.integers(20, 44))])
    return path
