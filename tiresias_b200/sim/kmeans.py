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
    """Returns (centroids, assignment, loss); ``assignment[i]`` is the queue of ``jobs[i]``.

    Same iteration as the reference (assign to the nearest centroid, first minimum wins; move each
    centroid to the member nearest the truncated mean; re-seed empty clusters at random; stop when
    the assignment repeats or after ``max_iter`` rounds), computed on a feature matrix. The
    iteration often falls into a cycle and the reference then burns all 1000 rounds; because a
    round without a random re-seed is a pure function of the centroids, a repeated centroid tuple
    lets us jump straight to the state round ``max_iter`` would reach — same result, ~100x less work.
    """
    if dist_fn is not None:
        return _clusterize_generic(jobs, k, dist_fn, max_iter, rng)
    rng = rng if rng is not None else np.random.mtrand._rand
    n = len(jobs)
    feats = np.array([job_features(j) for j in jobs], dtype=np.float64).reshape(n, 7)
    sums = np.array([transform_to_dist(j) for j in jobs], dtype=np.float64)

    def distances(cent: List[int]) -> np.ndarray:
        diff = np.abs(feats[:, None, :] - feats[cent][None, :, :])
        d = diff[:, :, 0].copy()
        for f in range(1, 7):                      # left-to-right adds, as the scalar code does
            d += diff[:, :, f]
        return d

    centroids = [int(i) for i in rng.randint(n, size=k)]
    new_assignment = None
    it = 0
    seen = {}                                      # centroid tuple -> round index
    history = {}                                   # round index -> (centroids, assignment)
    while it < max_iter:
        old_assignment = new_assignment
        it += 1
        new_assignment = np.argmin(distances(centroids), axis=1)
        reseeded = False
        for c in range(k):
            members = np.flatnonzero(new_assignment == c)
            if members.size:
                member_sums = sums[members]
                mean_score = int(np.mean(member_sums, axis=0))
                centroids[c] = int(members[int(np.argmin(np.abs(member_sums - mean_score)))])
            else:
                centroids[c] = int(rng.choice(n))
                reseeded = True
        if old_assignment is not None and np.array_equal(new_assignment, old_assignment):
            break
        if reseeded:
            seen.clear()
            history.clear()
            continue
        key = tuple(centroids)
        if key in seen:
            first = seen[key]
            period = it - first
            target = first + 1 + ((max_iter - first - 1) % period)
            centroids, new_assignment = history[target] if target != it else (centroids, new_assignment)
            centroids = list(centroids)
            break
        seen[key] = it
        history[it] = (tuple(centroids), new_assignment)
    assignment = [int(a) for a in new_assignment]
    final = distances(centroids)
    loss = 0
    for j in range(n):
        loss += float(final[j, assignment[j]])
    return [jobs[c] for c in centroids], assignment, loss


def _clusterize_generic(jobs, k, dist_fn, max_iter, rng):
    """Scalar version for a user-supplied distance function."""
    rng = rng if rng is not None else np.random.mtrand._rand
    n = len(jobs)
    sums = [transform_to_dist(j) for j in jobs]
    centroids = [int(i) for i in rng.randint(n, size=k)]
    new_assignment: List = [None] * n
    old_assignment: List = [None] * n
    it = 0
    while it < max_iter and (new_assignment != old_assignment or it == 0):
        old_assignment = list(new_assignment)
        it += 1
        for j in range(n):
            ds = [dist_fn(jobs[j], jobs[c]) for c in centroids]
            new_assignment[j] = int(np.argmin(ds))
        for c in range(k):
            members = [j for j in range(n) if new_assignment[j] == c]
            if members:
                mean_score = int(np.mean([sums[j] for j in members], axis=0))
                centroids[c] = min(members, key=lambda j: abs(sums[j] - mean_score))
            else:
                centroids[c] = int(rng.choice(n))
    loss = 0
    for j in range(n):
        loss += dist_fn(jobs[j], jobs[centroids[new_assignment[j]]])
    return [jobs[c] for c in centroids], new_assignment, loss
