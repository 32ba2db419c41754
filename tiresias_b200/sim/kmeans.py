"""Queue assignment for horus+: a k-medoid-style clustering of queued jobs (reference:
core/jobs/utils.py:4-68, C8).

Centroids are jobs. Distance = L1 over seven job features; a cluster's new centroid is the member
whose *scalarised* feature sum is closest to the (integer-truncated) mean of the members' sums.
Random initialisation and empty-cluster re-seeding come from the legacy NumPy generator, in the
reference's call order (``randint(len(jobs), size=k)``, then ``choice(len(jobs))``), so seeded runs
reproduce its queue layouts.

The feature rows are precomputed once per call (the reference recomputes every pairwise distance
from attributes); the additions keep the reference's left-to-right order so float sums are identical.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np


def job_features(job) -> Tuple[float, ...]:
    return (len(job.tasks), job.gpu_utilization_avg, job.gpu_per_worker, job.gpus,
            job.gpu_utilization_max, job.gpu_mem_avg, job.gpu_mem_max)


def job_dist(x, y):
    """L1 distance between two jobs over :func:`job_features`."""
    return _l1(job_features(x), job_features(y))


def _l1(a: Sequence[float], b: Sequence[float]):
    score = abs(a[0] - b[0])
    for i in range(1, 7):
        score += abs(a[i] - b[i])
    return score


def transform_to_dist(job):
    f = job_features(job)
    score = f[0]
    for i in range(1, 7):
        score += f[i]
    return score


def get_closest(jobs, score):
    best, best_gap = None, 99999999999
    for j in jobs:
        gap = abs(transform_to_dist(j) - score)
        if gap < best_gap:
            best, best_gap = j, gap
    assert best is not None
    return best


def clusterize(jobs: List, k: int, dist_fn=None, max_iter: int = 1000, rng=None):
    """Returns (centroids, assignment, loss); ``assignment[i]`` is the queue of ``jobs[i]``."""
    rng = rng if rng is not None else np.random.mtrand._rand
    n = len(jobs)
    feats = [job_features(j) for j in jobs]
    sums = [transform_to_dist(j) for j in jobs]
    index_of = {id(j): i for i, j in enumerate(jobs)}
    dist = (lambda a, b: _l1(feats[a], feats[b])) if dist_fn is None else \
        (lambda a, b: dist_fn(jobs[a], jobs[b]))

    centroids = [int(i) for i in rng.randint(n, size=k)]
    new_assignment: List = [None] * n
    old_assignment: List = [None] * n
    it = 0
    while it < max_iter and (new_assignment != old_assignment or it == 0):
        old_assignment = list(new_assignment)
        it += 1
        for j in range(n):
            best_c, best_d = 0, None
            for c in range(k):
                d = dist(j, centroids[c])
                if best_d is None or d < best_d:          # first minimum wins, like np.argmin
                    best_c, best_d = c, d
            new_assignment[j] = best_c
        for c in range(k):
            members = [j for j in range(n) if new_assignment[j] == c]
            if members:
                mean_score = int(np.mean([sums[j] for j in members], axis=0))
                pick, gap = None, 99999999999
                for j in members:
                    g = abs(sums[j] - mean_score)
                    if g < gap:
                        pick, gap = j, g
                centroids[c] = pick
            else:
                centroids[c] = int(rng.choice(n))
    loss = 0
    for j in range(n):
        loss += dist(j, centroids[new_assignment[j]])
    del index_of
    return [jobs[c] for c in centroids], new_assignment, loss
