"""Command-line flag registry with the reference's surface (core/flags.py:14-125, C1).

Behaviour kept: flags are declared up-front, parsed lazily on first attribute read with
``parse_known_args`` (unknown flags are ignored), boolean flags accept ``--x``,
``--x=True``, ``--x True`` and ``--nox``, values may be overridden by assignment, and ``-v`` /
``--version`` prints the version. The implementation is a small registry class instead of the
module-global parser of the reference, so several independent flag sets can coexist in-process
(tests, the batch driver running simulations without subprocesses).
"""
from __future__ import annotations

import argparse
from typing import Any, Dict, List, Optional, Sequence

_TRUE = ("true", "t", "1")


def _to_bool(text: str) -> bool:
    return text.lower() in _TRUE


class FlagSet:
    """A set of typed flags; attribute access triggers parsing of ``argv`` exactly once."""

    def __init__(self, description: str = "") -> None:
        object.__setattr__(self, "_parser", argparse.ArgumentParser(description=description))
        object.__setattr__(self, "_values", {})
        object.__setattr__(self, "_parsed", False)
        object.__setattr__(self, "_argv", None)
        object.__setattr__(self, "unparsed", [])

    # -- declaration ------------------------------------------------------------------------
    def _define(self, name: str, default: Any, doc: str, kind) -> None:
        self._parser.add_argument("--" + name, default=default, help=doc, type=kind)

    def define_string(self, name: str, default: Optional[str], doc: str = "") -> None:
        self._define(name, default, doc, str)

    def define_integer(self, name: str, default: Optional[int], doc: str = "") -> None:
        self._define(name, default, doc, int)

    def define_float(self, name: str, default: Optional[float], doc: str = "") -> None:
        self._define(name, default, doc, float)

    def define_boolean(self, name: str, default: bool, doc: str = "") -> None:
        self._parser.add_argument("--" + name, nargs="?", const=True, default=default, help=doc,
                                  type=_to_bool)
        self._parser.add_argument("--no" + name, action="store_false", dest=name.replace("-", "_"))

    define_bool = define_boolean

    def define_version(self, version: str) -> None:
        self._parser.add_argument("-v", "--version", action="version", version="%(prog)s " + version,
                                  help="display version information")

    # -- parsing ----------------------------------------------------------------------------
    def parse(self, argv: Optional[Sequence[str]] = None) -> List[str]:
        ns, rest = self._parser.parse_known_args(args=argv)
        self._values.update(vars(ns))
        object.__setattr__(self, "_parsed", True)
        object.__setattr__(self, "unparsed", rest)
        return rest

    def set_argv(self, argv: Optional[Sequence[str]]) -> None:
        """Arguments to use when the lazy parse fires (default: ``sys.argv[1:]``)."""
        object.__setattr__(self, "_argv", list(argv) if argv is not None else None)

    def __getattr__(self, name: str) -> Any:
        if name.startswith("_"):
            raise AttributeError(name)
        if not self._parsed:
            self.parse(self._argv)
        try:
            return self._values[name]
        except KeyError:
            raise AttributeError(name) from None

    def __setattr__(self, name: str, value: Any) -> None:
        if not self._parsed:
            self.parse(self._argv)
        self._values[name] = value

    def as_dict(self) -> Dict[str, Any]:
        if not self._parsed:
            self.parse(self._argv)
        return dict(self._values)


def define_simulator_flags(fs: Optional[FlagSet] = None) -> FlagSet:
    """The 18 flags + ``-v`` of the reference CLI with identical names and defaults
    (run_sim.py:19-93), plus the new-framework flags (all optional, all off by default)."""
    import time

    fs = fs or FlagSet("tiresias_b200 cluster scheduling simulator")
    fs.define_string("trace_file", "tf_job.csv", "job trace (*.csv ',' delimited, *.txt ' ' delimited)")
    fs.define_string("log_path", "result-" + time.strftime("%Y%m%d-%H-%M-%S", time.localtime()),
                     "simulation output folder under log/")
    fs.define_string("scheme", "yarn", "placement: count | yarn | random | crandom | greedy | balance | "
                                       "cbalance | horus | horus+ | gandiva")
    fs.define_string("schedule", "fifo", "scheduler: fifo | fjf | sjf | lpjf | shortest | shortest-gpu | "
                                         "shortest-expected | dlas | dlas-gpu | dlas-gpu-gittins | "
                                         "dlas-gpu-pack | multi-dlas-gpu | gittins | gandiva-legacy | "
                                         "gpu-demands | horus | horus+ | gandiva")
    fs.define_boolean("pack", False, "enable packing for gpu jobs")
    fs.define_integer("num_switch", 1, "number of switches (racks)")
    fs.define_integer("num_node_p_switch", 32, "nodes under one switch")
    fs.define_boolean("enable_network_costs", False, "lengthen cross-node jobs by the network cost model")
    fs.define_boolean("enable_migration", False, "enable preempt-and-migrate scan")
    fs.define_integer("bandwidth", 1250, "bandwidth per rack in MB/s (10 Gb/s)")
    fs.define_float("internode_latency", 0.015, "latency per crossed node in seconds")
    fs.define_integer("gpu_memory_capacity", 32, "GPU memory in GiB")
    fs.define_integer("num_queue", 1, "number of queues in the job manager")
    fs.define_integer("num_gpu_p_node", 8, "gpus per node")
    fs.define_integer("num_cpu_p_node", 128, "cpus per node")
    fs.define_integer("mem_p_node", 512, "memory per node")
    fs.define_string("cluster_spec", None, "cluster spec csv: num_switch,num_node_p_switch,"
                                           "num_gpu_p_node,num_cpu_p_node,mem_p_node")
    fs.define_boolean("print", False, "print debug information")
    fs.define_boolean("flush_stdout", True, "flush stdout")
    fs.define_version("0.1")
    # ---- new-framework flags ---------------------------------------------------------------
    fs.define_integer("seed", None, "seed numpy/python RNGs before the run (golden-log parity)")
    fs.define_string("mode", "parity", "parity: reproduce the reference bit for bit (inert "
                                       "interference/migration/network); calibrated: use measured "
                                       "B200 cost tables")
    fs.define_string("queue_limit", None, "MLFQ thresholds in seconds, comma separated "
                                          "(README: '3600' or '3600,7200,18000')")
    fs.define_string("job_dist", None, "duration samples for the Gittins index (csv with a "
                                       "'duration' column); default: the trace itself")
    fs.define_integer("gittins_delta", 3250, "Gittins service quantum")
    fs.define_boolean("gittins_desc", False, "gittins: highest index first (the reference's loop sorts "
                                             "ascending, its dlas-gpu-gittins variant descending)")
    fs.define_boolean("gittins_clean", False, "gittins without three accidents of the reference loop: waiting "
                                              "jobs ranked by attained GPU-time, a wake-up every gittins_delta, "
                                              "running-first / submit-order tie break (what exec/ executes)")
    fs.define_string("pack_mem_util", "unit", "GPU share a job takes under the packing schedulers (dlas-gpu-pack, "
                                              "gandiva-legacy): unit = 1 per job as in the reference's live table "
                                              "(packing never engages) | table = the reference's commented-out "
                                              "per-model shares (core/models.py:21)")
    fs.define_integer("pack_max_share", 3, "fifo-pack: single-GPU jobs per GPU (reference device limit: 3)")
    fs.define_string("pack_slowdown", "1,1.2,1.2", "fifo-pack: slowdown of a job sharing its GPU with 0, 1, 2 others "
                                                    "(reference: 1 + interference FACTOR 0.2; measured table from "
                                                    "calib/interference.py in calibrated runs)")
    fs.define_integer("slice_quantum", 100, "gandiva-slice: time slice in trace time units (reference: 100 ticks)")
    fs.define_float("solve_starvation", 0, "DLAS starvation knob: promote when pending >= x * executed")
    fs.define_integer("resume_overhead", 0, "time units charged to a job each time it resumes after a "
                                            "preemption (measured restart cost; 0 = free, as in the reference)")
    fs.define_string("calibration", None, "calibration json produced by tiresias_b200.calib")
    fs.define_string("comm", "ps", "communication mode priced/executed for distributed jobs: ps | ring")
    fs.define_boolean("quiet", False, "do not mirror the log to stderr")
    fs.define_boolean("fast", False, "skip the per-task Gaussian utilisation draws (every task reports its "
                                     "mean): no random stream is consumed, cluster.csv equals the exact "
                                     "mode's in every column but avg_gpu_utilization whenever the policy "
                                     "does not score devices (yarn placement)")
    return fs
