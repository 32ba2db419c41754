"""Priority math of the Tiresias schedulers, in PyTorch (BASELINE.json: "Scheduler priority math
(LAS, Gittins index) stays in PyTorch").

* Gittins index table (reference: run_sim.py:948-953 ``get_gittins_index``, :1649-1680
  ``cal_r_gittins_index``, :1682-1707 ``parse_job_dist``). For attained service ``a`` over the sorted
  duration samples ``data[0..num-1]`` and service quantum ``delta``::

      idx       = first i with data[i] > a
      idx_delta = first i with data[i] > a + delta      (num-1 if none)
      p   = round((idx_delta - idx) / (num - idx), 5)                   # P[finish within delta]
      e   = round((sum(data[idx:idx_delta]) + delta*(num - idx_delta)) / (num - idx), 5)
      r   = round(p * 1e6 / e, 4)

  The table holds ``r`` for ``a = data[i] - 1`` per sample, then a ``sys.maxsize`` sentinel with
  index 0. ``gittins_table`` computes the whole table at once with ``searchsorted`` + ``cumsum``
  (the reference does a linear scan and a slice sum per sample: O(n^2)); ``cal_r_gittins_index`` is
  the scalar restatement used as the test oracle.
* Discretised LAS: ``las_queue_after`` demotes a batch of jobs by attained service against the
  MLFQ thresholds.
"""
from __future__ import annotations

import csv
import sys
from typing import Dict, Optional, Sequence



class _LazyTorch:
    """``torch`` on first use: the priority math (LAS queues, Gittins ranks) is in PyTorch, but importing it
    costs seconds that fifo / sjf / shortest runs never need."""

    def __getattr__(self, name):
        import torch as _torch
        globals()["torch"] = _torch
        return getattr(_torch, name)


torch = _LazyTorch()

MAXSIZE = sys.maxsize


def _round_half_even(x: torch.Tensor, digits: int) -> torch.Tensor:
    """Python ``round(x, digits)`` on float64 tensors. Python rounds the exact decimal expansion;
    scaling by 10**digits and rounding half-to-even agrees except on exact binary ties, which the
    index's quotients do not hit in practice — the table test cross-checks against ``round``."""
    scale = 10.0 ** digits
    return torch.round(x * scale) / scale


def cal_r_gittins_index(job_data: Dict, a: int, delta: int) -> float:
    """Scalar oracle (one attained-service value)."""
    data, num = job_data["data"], job_data["num"]
    if a > data[num - 1] - 1:
        return 0.0
    idx = next(i for i, v in enumerate(data) if v > a)
    nxt = a + delta
    if nxt > data[num - 1] - 1:
        idx_delta = num - 1
    else:
        idx_delta = next(i for i, v in enumerate(data) if v > nxt)
    p = round(((idx_delta - idx) * 1.0) / (num - idx), 5)
    e_sum = sum(data[idx:idx_delta]) + delta * (num - idx_delta)
    e = round(e_sum / (num - idx), 5)
    return round(p * 1000000 / e, 4)


def gittins_table(durations: Sequence[int], delta: int) -> Dict:
    """{num, data (sorted + sentinel), gittins (per sample + 0.0)} — vectorised."""
    data = torch.tensor(sorted(int(d) for d in durations), dtype=torch.int64)
    num = int(data.numel())
    if num == 0:
        raise ValueError("no duration samples for the Gittins index")
    a = data - 1
    last = data[-1] - 1
    idx = torch.searchsorted(data, a, right=True)                  # first data > a
    nxt = a + delta
    idx_delta = torch.searchsorted(data, nxt, right=True)
    idx_delta = torch.where(nxt > last, torch.full_like(idx_delta, num - 1), idx_delta)
    prefix = torch.cat([torch.zeros(1, dtype=torch.int64), torch.cumsum(data, 0)])
    seg = prefix[idx_delta] - prefix[idx]
    seg = torch.where(idx_delta > idx, seg, torch.zeros_like(seg))
    denom = (num - idx).to(torch.float64)
    p = _round_half_even((idx_delta - idx).to(torch.float64) / denom, 5)
    e = _round_half_even((seg + delta * (num - idx_delta)).to(torch.float64) / denom, 5)
    r = _round_half_even(p * 1000000 / e, 4)
    r = torch.where(a > last, torch.zeros_like(r), r)
    table = [float(v) for v in r.tolist()]
    # exact-tie guard: re-derive entries where torch and python rounding could disagree
    job_data = {"num": num, "data": data.tolist()}
    for i in range(num):
        ref = cal_r_gittins_index(job_data, int(a[i]), delta) if _near_tie(p[i], e[i]) else table[i]
        table[i] = ref
    return {"num": num, "data": data.tolist() + [MAXSIZE], "gittins": table + [0.0],
            "data_t": data, "gittins_t": torch.tensor(table + [0.0], dtype=torch.float64)}


def _near_tie(p: torch.Tensor, e: torch.Tensor) -> bool:
    # the only fragile spots are values within 1e-9 of a rounding boundary; recompute those in python
    for v, d in ((float(p), 5), (float(e), 5)):
        frac = (v * 10 ** d) % 1.0
        if abs(frac - 0.5) < 1e-6:
            return True
    return False


def get_gittins_index(job_dist: Dict, a) -> float:
    """Rank of a job with attained service ``a`` (0 beyond the second-largest sample)."""
    data = job_dist["data"]
    if a > data[-2]:
        return 0
    t = job_dist.get("data_t")
    if t is not None:
        idx = int(torch.searchsorted(t, torch.tensor(int(a)), right=True))
        idx = min(idx, len(data) - 1)
    else:
        idx = next(i for i, v in enumerate(data) if v > a)
    return job_dist["gittins"][idx]


def gittins_ranks(job_dist: Dict, attained: torch.Tensor) -> torch.Tensor:
    """Batched :func:`get_gittins_index` for a vector of attained-service values."""
    t, g = job_dist["data_t"], job_dist["gittins_t"]
    idx = torch.searchsorted(t, attained.to(torch.int64), right=True).clamp_(max=g.numel() - 1)
    ranks = g[idx]
    second_largest = job_dist["data"][-2]
    return torch.where(attained > second_largest, torch.zeros_like(ranks), ranks)


def parse_job_dist(path: Optional[str], delta: int, fallback_durations: Optional[Sequence[int]] = None) -> Dict:
    """Duration samples from a csv with a ``duration`` column (the reference hard-codes
    ``yarn-gput1000.csv``, which does not ship); without a file the trace's own durations are used."""
    if path:
        with open(path, "r") as fh:
            durations = [int(float(row["duration"])) for row in csv.DictReader(fh)]
    else:
        durations = list(fallback_durations or [])
    return gittins_table(durations, delta)


def cal_shortest_expected_remaining(job_data: Dict, a) -> float:
    """Mean of the duration samples above ``a`` (run_sim.py:288-296; the slice stops one short of
    the last real sample, as in the reference)."""
    data, num = job_data["data"], job_data["num"]
    idx = next(i for i, v in enumerate(data) if v > a)
    if idx == num - 1:
        return data[idx]
    return round(sum(data[idx:num - 1]) * 1.0 / (num - 1 - idx), 2)


def las_queue_after(attained: torch.Tensor, q_id: torch.Tensor, limits: Sequence[int]) -> torch.Tensor:
    """One demotion step of discretised LAS: a job in queue q < last moves to q+1 when its attained
    service reached ``limits[q]`` (run_sim.py:746-758 demotes one level per event)."""
    lim = torch.tensor(list(limits) + [MAXSIZE], dtype=torch.int64)
    last = len(limits)
    demote = (q_id < last) & (attained.to(torch.int64) >= lim[q_id.clamp(max=last)])
    return q_id + demote.to(q_id.dtype)
