"""``python run_sim.py ...`` — one simulation run (reference: run_sim.py:19-96, :1709-1756; C2).

Same flags and defaults as the reference, same output layout
``log/<log_path>/<%Y-%m-%d-%H-%M-%S-%f>/{output.log,cluster.csv,job.csv,cpu.csv,gpu.csv,memory.csv,
network.csv}`` plus ``jobs.csv`` (per-job rows under the README's file name) and ``summary.json``.

``--schedule`` selects the engine: fifo / horus / horus+ / gandiva run on the tick engine
(:mod:`tiresias_b200.sim.scheduler`), the Tiresias family (sjf, lpjf, shortest, shortest-gpu, dlas,
dlas-gpu, gittins, ...) on the event engine (:mod:`tiresias_b200.sim.legacy`). A live-schema trace
given to an event scheduler (or the reverse) is converted on the fly.
"""
from __future__ import annotations

import datetime
import json
import logging
import os
import random
from typing import Dict, List, Optional, Sequence

import numpy as np

from ..sim import util
from ..sim.flags import FlagSet, define_simulator_flags

TICK_SCHEDULES = ("fifo", "horus", "horus+", "gandiva")


def _attach_logging(output_dir: str, quiet: bool) -> List[logging.Handler]:
    log = util.logger
    log.setLevel(logging.DEBUG)
    log.propagate = False
    handlers: List[logging.Handler] = []
    fh = logging.FileHandler(os.path.join(output_dir, "output.log"), mode="w")
    fh.setFormatter(logging.Formatter("%(asctime)s %(levelname)s: %(message)s"))
    handlers.append(fh)
    if not quiet:
        sh = logging.StreamHandler()
        sh.setFormatter(logging.Formatter(
            "%(asctime)s p%(process)s {%(module)s:%(lineno)d} %(levelname)s: %(message)s"))
        handlers.append(sh)
    for h in handlers:
        log.addHandler(h)
    return handlers


def load_calibration(path: Optional[str]) -> Optional[Dict]:
    if not path:
        return None
    with open(path, "r") as fh:
        return json.load(fh)


def run(flags: FlagSet, output_dir: Optional[str] = None) -> Dict:
    """Run one simulation with already-parsed ``flags``; returns the summary dict (also written to
    ``summary.json``). ``output_dir`` overrides the ``log/<log_path>/<timestamp>`` default."""
    if output_dir is None:
        stamp = datetime.datetime.now().strftime("%Y-%m-%d-%H-%M-%S-%f")
        output_dir = os.path.join("log", flags.log_path, stamp)
    os.makedirs(output_dir, exist_ok=True)
    handlers = _attach_logging(output_dir, bool(flags.quiet))
    try:
        rng = None
        if flags.seed is not None:
            np.random.seed(flags.seed)
            random.seed(flags.seed)
        calibration = load_calibration(flags.calibration)
        if getattr(flags, "fast", False):
            from ..sim.infra import MeanRandomState, global_rng
            rng = MeanRandomState(global_rng())
        if _use_tick_engine(flags):
            summary = _run_tick(flags, output_dir, rng, calibration)
        else:
            from ..sim.legacy.engine import run_event_simulation
            summary = run_event_simulation(flags, output_dir, calibration)
        summary.update({"scheme": flags.scheme, "schedule": flags.schedule,
                        "num_queue": flags.num_queue, "trace_file": flags.trace_file,
                        "output_dir": output_dir})
        with open(os.path.join(output_dir, "summary.json"), "w") as fh:
            json.dump(summary, fh, indent=1, default=float)
        return summary
    finally:
        for h in handlers:
            util.logger.removeHandler(h)
            h.close()


def _use_tick_engine(flags) -> bool:
    """horus / horus+ / gandiva always tick; fifo ticks on a live-schema trace (the reference's live
    path) and runs on the event engine (Tiresias ``sim_job_events``) for a README-schema trace or
    ``--scheme count``."""
    from ..sim.trace import detect_schema
    if flags.schedule not in TICK_SCHEDULES or flags.scheme == "count":
        return False
    if flags.schedule == "fifo" and flags.trace_file and os.path.exists(flags.trace_file):
        return detect_schema(flags.trace_file) == "live"
    return True


def _run_tick(flags: FlagSet, output_dir: str, rng, calibration) -> Dict:
    from ..sim.infra import Infrastructure
    from ..sim.jobs_manager import JobsManager
    from ..sim.logs import LogManager
    from ..sim.queues import JobQueueManager
    from ..sim.scheduler import Scheduler
    from ..sim.trace import detect_schema

    log_manager = LogManager(output_dir, flags)
    infrastructure = Infrastructure(flags, rng=rng)
    log_manager.init(infrastructure)
    reader = None
    if flags.trace_file and os.path.exists(flags.trace_file) and detect_schema(flags.trace_file) == "tiresias":
        from ..sim.legacy.engine import tiresias_trace_as_live
        reader = tiresias_trace_as_live(flags.trace_file)
    jq = JobQueueManager(flags, flags.trace_file)
    jm = JobsManager(flags, jq, rng=rng, trace_reader=reader)
    if calibration and calibration.get("interference_factor") is not None and flags.mode == "calibrated":
        jm.set_interference_factor(float(calibration["interference_factor"]), infrastructure)
    if calibration and flags.mode == "calibrated":
        jm.set_measured_profiles(calibration)            # K7: measured utilisation / memory per model
    sched = Scheduler(infrastructure, jm, log_manager, enable_migration=flags.enable_migration,
                      calibration=calibration)
    try:
        sched.start()
    finally:
        log_manager.close()
    return sched.summary()


def main(argv: Optional[Sequence[str]] = None) -> Dict:
    flags = define_simulator_flags()
    flags.parse(list(argv) if argv is not None else None)
    return run(flags)


if __name__ == "__main__":
    main()
