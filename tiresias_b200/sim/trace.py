"""Job sources (reference: core/jobs/job_generator.py:14-207, C10c; README-schema parser
core/jobs/job_queue_manager.py:29-62, D15).

* :class:`JobTraceReader` — the live pandas schema (``type, normalized_time, minutes,
  gpu_per_container, used_gpus, gpu_utilization_avg, gpu_utilization_max, memory_max,
  memory_avg``). Preparation is the reference's, step for step: keep
  ``type == noninteractive``, sort by ``normalized_time``, drop rows with any NaN, shift to zero,
  divide by 10000. Arrival lookup, however, is a cursor over the sorted rows instead of two
  full-frame boolean masks per tick (the reference's hot loop (a)).
* :class:`JobGenerator` — empirical-CDF sampler used by ``demonstrate_trace.py``.
* :func:`read_tiresias_trace` — the README schema ``job_id,num_gpu,submit_time,iterations,
  model_name,duration,interval`` (csv ``,`` or txt space-delimited) for the event-driven schedulers.
"""
from __future__ import annotations

import csv
import logging
import os
import random
import sys
from typing import Dict, List, Optional, Sequence

import numpy as np

from .util import logger

LIVE_COLUMNS = ("type", "normalized_time", "minutes", "gpu_per_container", "used_gpus",
                "gpu_utilization_avg", "gpu_utilization_max", "memory_max", "memory_avg")
README_COLUMNS = ("job_id", "num_gpu", "submit_time", "iterations", "model_name", "duration", "interval")
TIME_DIVISOR = 10000          # core/jobs/job_generator.py:189


def detect_schema(path: str) -> str:
    """"live" (pandas schema) or "tiresias" (README schema), from the header line."""
    with open(path, "r") as fh:
        header = fh.readline()
    names = [c.strip() for c in header.replace(",", " ").split()]
    if "normalized_time" in names and "used_gpus" in names:
        return "live"
    if "num_gpu" in names and "submit_time" in names:
        return "tiresias"
    raise ValueError(f"{path}: unknown trace schema (header: {header.strip()!r})")


class TraceRow:
    """One arrival of the live schema; attribute names are the csv columns."""

    __slots__ = ("index", "normalized_time", "minutes", "gpu_per_container", "used_gpus",
                 "gpu_utilization_avg", "gpu_utilization_max", "memory_max", "memory_avg", "model_name",
                 "iterations")

    def __init__(self, index, **kw):
        self.index = index
        self.model_name = kw.pop("model_name", None)
        self.iterations = kw.pop("iterations", None)
        for k, v in kw.items():
            setattr(self, k, v)


class JobTraceReader:
    """Replays a live-schema trace in arrival order."""

    def __init__(self, file_path: str) -> None:
        import pandas as pd
        if not os.path.exists(file_path):
            logging.error(f"file: {file_path} not exist")
            sys.exit(1)
        try:
            self.trace_df = pd.read_csv(file_path)
        except Exception:
            logging.error(f"unable to read the file, assumed it was csv. but got {file_path}")
            sys.exit(1)
        self.delta_time_offset = 0
        self._cursor = 0
        self._rows: List[TraceRow] = []
        self._times = np.zeros(0)

    def prepare_jobs(self) -> None:
        df = self.trace_df
        logger.info("original: %d" % len(df))
        df = df[df["type"] == "noninteractive"]
        logger.info("only batch: %d" % len(df))
        df = df.sort_values(by="normalized_time")
        df = df.dropna()
        logger.info("dropped nan: %d" % len(df))
        df = df.copy()
        df["normalized_time"] -= df["normalized_time"].min()
        df["normalized_time"] /= TIME_DIVISOR
        logger.info("Finished Prepping...")
        self.trace_df = df
        cols = {c: df[c].to_numpy() for c in LIVE_COLUMNS[1:]}
        extra = {c: df[c].to_numpy() for c in ("model_name", "iterations") if c in df.columns}
        labels = df.index.to_numpy()
        self._rows = []
        for i in range(len(df)):
            kw = {c: cols[c][i].item() if hasattr(cols[c][i], "item") else cols[c][i] for c in cols}
            for c in extra:
                v = extra[c][i]
                kw[c] = v.item() if hasattr(v, "item") else v
            lab = labels[i]
            self._rows.append(TraceRow(lab.item() if hasattr(lab, "item") else lab, **kw))
        self._times = np.asarray(cols["normalized_time"], dtype=np.float64)
        self._cursor = 0

    def __len__(self) -> int:
        return len(self._rows)

    def remaining_jobs(self) -> int:
        return len(self._rows) - self._cursor

    def generate_jobs(self, delta_time) -> List[TraceRow]:
        """Rows that arrived by ``delta_time`` and were not handed out yet (arrival order)."""
        end = int(np.searchsorted(self._times, delta_time, side="right"))
        if end <= self._cursor:
            return []
        out = self._rows[self._cursor:end]
        self._cursor = end
        return out

    def all_rows(self) -> List[TraceRow]:
        return list(self._rows)


class JobGenerator:
    """Samples synthetic traces from empirical populations. ``cdf`` gives the evenly spaced
    cumulative weights the reference uses (it does not depend on the sample values). The model
    "distribution" defaults to the raw model samples, as in the reference
    (core/jobs/job_generator.py:46), so model sizes are drawn with those values as cumulative
    weights."""

    MODEL_SAMPLE_DEF = [1.04, 1.27, 15.36, 108.48, 125.16, 125.16, 237.61, 330.52, 364.2, 462.77,
                        544.24, 636.75, 664.1, 700., 798.28, 862.95]
    ITTER_SAMPLE_DEF = [1, 1, 1, 1, 1, 1, 109, 126, 133, 138, 141, 143, 144, 147, 157, 168, 175, 192,
                        193, 198, 235, 237, 242, 253, 258, 272, 272, 274, 288, 326, 326, 362, 386, 391,
                        410, 438, 447, 468, 473, 513, 513, 521, 521, 525, 581, 606, 607, 775, 775, 789,
                        822, 864, 864, 892, 903, 949, 1011, 1085, 1360, 1501, 2178, 2239, 2275, 3304,
                        3469, 4861]
    DURATION_SAMPLE_DEF = [121] * 21 + [122] * 4 + [123] * 3 + [124, 125, 125, 126, 126, 126, 126, 127,
                                                              128, 130, 131, 133, 135, 138, 141, 143, 147,
                                                              152, 155, 158, 164, 171, 180, 189, 196, 209,
                                                              230, 263, 305, 368, 536, 1800]
    ARRIVAL_SAMPLE_DEF = [10, 100, 500, 1000, 3000, 5000, 10000]

    def __init__(self) -> None:
        self.MODEL_DIST_DEF = self.cdf(self.MODEL_SAMPLE_DEF)
        self.ITTER_DIST_DEF = self.cdf(self.ITTER_SAMPLE_DEF)
        self.ARRIVAL_DIST_DEF = self.cdf(self.ARRIVAL_SAMPLE_DEF)
        self.DURATION_DIST_DEF = self.cdf(self.DURATION_SAMPLE_DEF)
        self.model_distribution = self.MODEL_SAMPLE_DEF
        self.model_samples = self.MODEL_SAMPLE_DEF
        self.duration_distribution = self.DURATION_DIST_DEF
        self.duration_samples = self.DURATION_SAMPLE_DEF
        self.itter_distribution = self.ITTER_DIST_DEF
        self.itter_samples = self.ITTER_SAMPLE_DEF
        self.arrival_distribution = self.ARRIVAL_DIST_DEF
        self.arrival_samples = self.ARRIVAL_SAMPLE_DEF

    @staticmethod
    def cdf(samples: Sequence[float]) -> np.ndarray:
        n = len(samples)
        return 1.0 * np.arange(n) / (n - 1)

    # accessor pairs kept for drop-in use
    def get_model_distribution(self): return self.model_distribution
    def get_duration_distribution(self): return self.duration_distribution
    def get_itter_distribution(self): return self.itter_distribution
    def get_arrival_distribution(self): return self.arrival_distribution
    def get_model_samples(self): return self.model_samples
    def get_duration_samples(self): return self.duration_samples
    def get_itter_samples(self): return self.itter_samples
    def get_arrival_samples(self): return self.arrival_samples

    def _set(self, what: str, samples) -> None:
        setattr(self, what + "_samples", samples)
        setattr(self, what + "_distribution", self.cdf(samples))

    def set_model_distribution(self, samples): self._set("model", samples)
    def set_duration_distribution(self, samples): self._set("duration", samples)
    def set_itter_distribution(self, samples): self._set("itter", samples)
    def set_arrival_distribution(self, samples): self._set("arrival", samples)

    @staticmethod
    def generate_samples(number: int, population, distribution) -> list:
        if not any(population) or not any(distribution):
            raise Exception("Error population and samples cannot be null")
        if number < len(population):
            print("WARN: generating fewer samples then population")
        return random.choices(population, cum_weights=distribution, k=number)

    def generate_trace(self, number: int) -> Dict[str, list]:
        """Columns ``model, duration, itter, arrival, num_gpu, interval, job_id`` (same draw order
        as the reference, so a seeded ``random`` gives the same trace)."""
        trace = {
            "model": self.generate_samples(number, self.model_samples, self.model_distribution),
            "duration": self.generate_samples(number, self.duration_samples, self.duration_distribution),
            "itter": self.generate_samples(number, self.itter_samples, self.itter_distribution),
            "arrival": self.generate_samples(number, self.arrival_samples, self.arrival_distribution),
        }
        trace["num_gpu"] = [random.randrange(1, 128) for _ in range(number)]
        trace["interval"] = [random.randrange(20, 44) for _ in range(number)]
        trace["job_id"] = list(range(number))
        return trace


# ---- README schema ---------------------------------------------------------------------------------

def _int_or_float(text: str):
    v = float(text)
    return int(v) if v == int(v) else v


def read_tiresias_trace(path: str) -> List[dict]:
    """Rows of a README-schema trace as dicts with typed values. ``*.csv`` is comma-delimited,
    ``*.txt`` space-delimited (README.md:17-25)."""
    if not os.path.exists(path):
        raise ValueError(f"trace file {path} does not exist")
    deli = " " if path.endswith(".txt") else ","
    rows: List[dict] = []
    with open(path, "r") as fh:
        reader = csv.DictReader(fh, delimiter=deli, skipinitialspace=True)
        for raw in reader:
            if raw.get("job_id") in (None, ""):
                continue
            row = {k.strip(): (v.strip() if isinstance(v, str) else v) for k, v in raw.items() if k}
            rows.append({
                "job_id": int(float(row["job_id"])),
                "num_gpu": int(float(row["num_gpu"])),
                "submit_time": _int_or_float(row["submit_time"]),
                "iterations": int(float(row.get("iterations") or 0)),
                "model_name": row.get("model_name") or "",
                "duration": int(float(row["duration"])),
                "interval": int(float(row.get("interval") or 0)),
            })
    return rows


def live_rows_as_tiresias(reader: JobTraceReader, scale_factor: float = 0.5,
                          default_model: Optional[str] = None) -> List[dict]:
    """View a live-schema trace through the README schema so every scheduler can replay it:
    ``duration = minutes * scale_factor`` (rounded up to >= 1), ``submit_time = int(normalized)``,
    ``num_gpu = used_gpus``; the model is the row's ``model_name`` or a size-class pick from the
    row's memory footprint."""
    from .models import model_for_memory
    out = []
    for r in reader.all_rows():
        dur = r.minutes * scale_factor
        out.append({
            "job_id": r.index, "num_gpu": int(r.used_gpus), "submit_time": int(r.normalized_time),
            "iterations": int(r.iterations) if r.iterations else 0,
            "model_name": r.model_name or default_model or model_for_memory(r.memory_max / (1 << 20)),
            "duration": max(1, int(-(-dur // 1))), "interval": 0,
        })
    return out
