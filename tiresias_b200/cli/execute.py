"""``python execute.py`` — batch driver (reference: execute.py:5-57, C3).

Runs ``run_sim.py`` once per (scheme, schedule, num_queue) x ``repeats`` as a subprocess — the only
process boundary the reference has — with the reference's hard-coded experiment: 4 switches x 32
nodes, trace ``data/<trace>.csv``, ``--enable_network_costs False --enable_migration True``, logs under
``log/nodes_p_s32_job_<trace>/<scheme>_<schedule>``. The reference is edited in place to change the
experiment (README.md:29-37); here the same knobs are also command-line options, and
``--in-process`` runs the simulations inside this interpreter (no 2-3 s import cost per run).
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
import time
from typing import List, Optional, Sequence, Tuple

DEFAULT_CONFIGS: List[Tuple[str, str, int]] = [("horus+", "horus+", 3), ("horus+", "horus+", 4),
                                               ("horus+", "horus+", 5)]       # execute.py:47-49
ALL_CONFIGS: List[Tuple[str, str, int]] = [("horus+", "horus+", 4), ("horus+", "horus+", 5),
                                           ("horus", "horus", 1), ("gandiva", "gandiva", 1),
                                           ("yarn", "fifo", 1)]               # execute.py:44-46


def build_argv(scheme: str, schedule: str, num_queue: int, trace_file: str = "month",
               num_nodes_p_switch: int = 32, num_switch: int = 4, migrate: bool = True,
               extra: Sequence[str] = ()) -> List[str]:
    trace_name = os.path.splitext(os.path.basename(trace_file))[0]
    log_sub_dir = "nodes_p_s" + str(num_nodes_p_switch) + "_job_" + trace_name
    trace_path = trace_file if os.path.exists(trace_file) else os.path.join("data", trace_file + ".csv")
    return ["--num_node_p_switch", str(num_nodes_p_switch), "--num_switch", str(num_switch),
            "--scheme", scheme, "--trace_file", trace_path, "--num_queue", str(num_queue),
            "--schedule", schedule, "--enable_network_costs", "False", "--enable_migration", str(migrate),
            "--log_path", os.path.join(log_sub_dir, f"{scheme}_{schedule}"), *extra]


def do_once(scheme: str, schedule: str, num_queue: int, in_process: bool = False, poll_s: float = 5.0,
            **kw) -> int:
    argv = build_argv(scheme, schedule, num_queue, **kw)
    if in_process:
        from .run_sim import main as run_main
        run_main(argv)
        return 0
    root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    proc = subprocess.Popen([sys.executable, os.path.join(root, "run_sim.py"), *argv])
    print("process pid %d: " % proc.pid)
    try:
        while proc.poll() is None:
            time.sleep(poll_s)
    except KeyboardInterrupt:
        proc.kill()
    return proc.returncode or 0


def main(argv: Optional[Sequence[str]] = None) -> None:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--trace", default="month", help="trace name under data/ (or a path)")
    ap.add_argument("--repeats", type=int, default=3)
    ap.add_argument("--num_switch", type=int, default=4)
    ap.add_argument("--num_node_p_switch", type=int, default=32)
    ap.add_argument("--all", action="store_true", help="also run horus, gandiva and yarn/fifo")
    ap.add_argument("--config", action="append", default=[], metavar="SCHEME:SCHEDULE:NUM_QUEUE")
    ap.add_argument("--in-process", action="store_true")
    ap.add_argument("--poll", type=float, default=5.0)
    args, extra = ap.parse_known_args(list(argv) if argv is not None else None)
    configs = [tuple(c.split(":")) for c in args.config] or (ALL_CONFIGS if args.all else DEFAULT_CONFIGS)
    for scheme, schedule, num_q in configs:
        for _ in range(args.repeats):
            do_once(scheme, schedule, int(num_q), in_process=args.in_process, poll_s=args.poll,
                    trace_file=args.trace, num_nodes_p_switch=args.num_node_p_switch,
                    num_switch=args.num_switch, extra=extra)


if __name__ == "__main__":
    main()
