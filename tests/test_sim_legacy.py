"""Hand-computed micro-traces for the event-driven Tiresias schedulers.

The reference's implementations are dead code that cannot run (run_sim.py:98-1707), so the oracle
is the schedule worked out by hand from those loops' rules; each expectation below is derived in
the comment next to it. Cluster: 1 switch x 1 node x 4 GPUs unless stated.
"""
import csv
import os

import pytest

from tiresias_b200.cli.run_sim import run
from tiresias_b200.sim.flags import define_simulator_flags

HEADER = "job_id,num_gpu,submit_time,iterations,model_name,duration,interval\n"


def _sim(tmp_path, jobs, schedule, scheme="count", extra=(), gpus=4, nodes=1):
    trace = tmp_path / "trace.csv"
    trace.write_text(HEADER + "".join(f"{i},{g},{s},100,resnet50,{d},30\n" for i, (g, s, d) in enumerate(jobs)))
    flags = define_simulator_flags()
    flags.parse(["--trace_file", str(trace), "--schedule", schedule, "--scheme", scheme, "--num_switch", "1",
                 "--num_node_p_switch", str(nodes), "--num_gpu_p_node", str(gpus), "--quiet", *extra])
    out = str(tmp_path / "out")
    summary = run(flags, output_dir=out)
    with open(os.path.join(out, "jobs.csv")) as fh:
        rows = {int(r["job_id"]): r for r in csv.DictReader(fh) if "job_id" in r}
    return rows, summary, out


def _ints(row, *keys):
    return tuple(int(row[k]) for k in keys)


def test_fifo_event_loop_with_pending_skip(tmp_path):
    # j0 4G@0 x100, j1 2G@10 x50, j2 1G@20 x10 on 4 GPUs. j0 runs 0-100. j1, j2 pend; pending is
    # sorted by num_gpu -> [j2, j1]. At t=100 j2 starts (100-110); removing it while iterating the
    # pending list skips j1 (reference loop, run_sim.py:1566-1575), which starts at the next event,
    # t=110, and ends at 160.
    rows, s, _ = _sim(tmp_path, [(4, 0, 100), (2, 10, 50), (1, 20, 10)], "fjf")
    assert _ints(rows[0], "start_time", "end_time", "JCT") == (0, 100, 100)
    assert _ints(rows[2], "start_time", "end_time", "JCT", "pending_time") == (100, 110, 90, 80)
    assert _ints(rows[1], "start_time", "end_time", "JCT", "pending_time") == (110, 160, 150, 100)
    assert s["makespan"] == 160 and s["finished_jobs"] == 3


def test_sjf_preempts_larger_job(tmp_path):
    # same trace, smallest-num_gpu-first with preemption: t=10 j1(2G) displaces j0(4G) (preempt 1);
    # t=20 j2 joins (3 GPUs busy), ends 30; j1 ends 60; j0 resumes 60 with 90 s left -> ends 150.
    rows, s, _ = _sim(tmp_path, [(4, 0, 100), (2, 10, 50), (1, 20, 10)], "sjf")
    assert _ints(rows[1], "start_time", "end_time", "JCT") == (10, 60, 50)
    assert _ints(rows[2], "start_time", "end_time", "JCT") == (20, 30, 10)
    assert _ints(rows[0], "start_time", "end_time", "JCT", "pending_time", "preempt", "resume") == \
        (0, 150, 150, 50, 1, 2)


def test_shortest_vs_sjf(tmp_path):
    # j0 2G@0 x100, j1 4G@10 x20. sjf keeps j0 (smaller) and runs j1 100-120; shortest-remaining
    # preempts j0 at t=10 (90 left vs 20), j1 runs 10-30, j0 resumes and ends at 120.
    jobs = [(2, 0, 100), (4, 10, 20)]
    rows, _, _ = _sim(tmp_path / "a", jobs, "sjf") if (tmp_path / "a").mkdir() is None else None
    assert _ints(rows[1], "start_time", "end_time") == (100, 120)
    assert _ints(rows[0], "end_time", "preempt") == (100, 0)
    (tmp_path / "b").mkdir()
    rows, _, _ = _sim(tmp_path / "b", jobs, "shortest")
    assert _ints(rows[1], "start_time", "end_time") == (10, 30)
    assert _ints(rows[0], "end_time", "preempt", "pending_time") == (120, 1, 20)


def test_shortest_gpu_orders_by_gpu_time(tmp_path):
    # j0 1G x100 (100 GPU-s), j1 4G x30 (120 GPU-s), both at t=0. shortest: j1 first (30 < 100),
    # j0 30-130. shortest-gpu: j0 first (100 < 120), j1 cannot fit beside it, runs 100-130.
    jobs = [(1, 0, 100), (4, 0, 30)]
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    rows, _, _ = _sim(tmp_path / "a", jobs, "shortest")
    assert _ints(rows[1], "start_time", "end_time") == (0, 30)
    assert _ints(rows[0], "start_time", "end_time") == (30, 130)
    rows, _, _ = _sim(tmp_path / "b", jobs, "shortest-gpu")
    assert _ints(rows[0], "start_time", "end_time") == (0, 100)
    assert _ints(rows[1], "start_time", "end_time") == (100, 130)


def test_dlas_demotes_at_queue_limit(tmp_path):
    # 2 queues, limit [50] (seconds of service). j0 4G@0 x100, j1 4G@10 x30.
    # t=50 (queue-jump wake-up): j0 attained 50 -> Q1; Q0's j1 takes the GPUs (start 50, end 80);
    # j0 preempted, resumes at 80 with 50 s left -> ends 130.
    rows, _, out = _sim(tmp_path, [(4, 0, 100), (4, 10, 30)], "dlas", extra=["--queue_limit", "50"])
    assert _ints(rows[1], "start_time", "end_time", "JCT", "pending_time") == (50, 80, 70, 40)
    assert _ints(rows[0], "start_time", "end_time", "JCT", "pending_time", "preempt", "resume", "promote") == \
        (0, 130, 130, 30, 1, 2, 0)
    times = [int(r["time"]) for r in csv.DictReader(open(os.path.join(out, "cluster.csv")))]
    assert times == [0, 10, 50, 80, 130]


def test_dlas_gpu_uses_gpu_time_and_reference_jump_formula(tmp_path):
    # limit [100] GPU-seconds. j0 2G@0 x100, j1 4G@5 x20. j0 crosses 100 GPU-s at t=50, but the
    # reference's wake-up is ceil((limit - executed)/num_gpu) + now = ceil(95/2)+5 = 53
    # (run_sim.py:929), so the demotion is noticed at t=53: j1 runs 53-73, j0 resumes, ends 120.
    rows, _, out = _sim(tmp_path, [(2, 0, 100), (4, 5, 20)], "dlas-gpu", extra=["--queue_limit", "100"])
    assert _ints(rows[1], "start_time", "end_time", "pending_time") == (53, 73, 48)
    assert _ints(rows[0], "end_time", "pending_time", "preempt") == (120, 20, 1)


def test_mlfq_four_queues_readme_limits(tmp_path):
    # README: 4 queues <-> queue_limit [3600, 7200, 18000]. A lone 1-GPU 20000 s job wakes the loop at
    # each threshold and finishes undisturbed.
    rows, _, out = _sim(tmp_path, [(1, 0, 20000)], "dlas", extra=["--queue_limit", "3600,7200,18000"])
    times = [int(r["time"]) for r in csv.DictReader(open(os.path.join(out, "cluster.csv")))]
    assert times == [0, 3600, 7200, 18000, 20000]
    assert _ints(rows[0], "end_time", "preempt", "promote") == (20000, 0, 0)


def test_dlas_starvation_promotes_back_to_q0(tmp_path):
    # limit [50], solve_starvation 1: j0 is demoted at 50 and preempted by j1 (4G@10 x200). Once j0
    # has waited as long as it has run since its last promotion (50 s) it returns to Q0 (promote 1)
    # — at t=100 — but Q0 is served in list order with running jobs in front, so it keeps waiting
    # until j1 itself is demoted at t=100 (attained 50): then Q0 = [j0], j0 runs 100-150.
    rows, _, _ = _sim(tmp_path, [(4, 0, 100), (4, 10, 200)], "dlas",
                      extra=["--queue_limit", "50", "--solve_starvation", "1"])
    assert int(rows[0]["promote"]) == 1
    assert _ints(rows[0], "end_time") == (150,)
    assert int(rows[1]["preempt"]) >= 1


def test_yarn_scheme_consolidates_per_switch(tmp_path):
    # 2 nodes x 4 GPUs: an 8-GPU job needs two completely idle nodes (MS-YARN); a 1-GPU job running
    # on node 0 blocks it until that job ends at t=50.
    rows, _, _ = _sim(tmp_path, [(1, 0, 50), (8, 10, 30)], "fjf", scheme="yarn", nodes=2)
    assert _ints(rows[1], "start_time", "end_time") == (50, 80)


def test_job_csv_row_layout(tmp_path):
    rows, _, out = _sim(tmp_path, [(4, 0, 100), (2, 10, 50)], "sjf")
    head = open(os.path.join(out, "job.csv")).readline().strip().split(",")
    assert head == ["time", "job_id", "num_gpu", "submit_time", "start_time", "end_time", "executed_time", "JCT",
                    "duration", "pending_time", "preempt", "resume", "promote"]      # --scheme count
    assert open(os.path.join(out, "job.csv")).read() == open(os.path.join(out, "jobs.csv")).read()
    (tmp_path / "y").mkdir()
    _, _, out = _sim(tmp_path / "y", [(4, 0, 100)], "sjf", scheme="yarn")
    head = open(os.path.join(out, "job.csv")).readline().strip().split(",")
    assert "resume" not in head and head[-2:] == ["preempt", "promote"]
    for f in ("cpu.csv", "gpu.csv", "network.csv", "memory.csv"):
        assert os.path.exists(os.path.join(out, f))


@pytest.mark.parametrize("schedule,scheme,extra", [
    ("lpjf", "yarn", ()), ("shortest-expected", "count", ()), ("dlas-gpu-gittins", "count", ("--queue_limit", "3600")),
    ("dlas-gpu-pack", "count", ()), ("multi-dlas-gpu", "yarn", ()), ("gittins", "count", ()),
    ("gittins", "count", ("--gittins_desc",)), ("gandiva-legacy", "gandiva", ()), ("sjf", "random", ()),
    ("sjf", "crandom", ()), ("sjf", "greedy", ()), ("sjf", "balance", ()), ("sjf", "cbalance", ()),
])
def test_every_schedule_and_scheme_completes_the_trace(tmp_path, schedule, scheme, extra):
    from tiresias_b200.sim.synth import tiresias_trace
    trace = tiresias_trace(str(tmp_path / "t.csv"), 40, seed=7, mean_gap=90)
    flags = define_simulator_flags()
    flags.parse(["--trace_file", trace, "--schedule", schedule, "--scheme", scheme, "--num_switch", "2",
                 "--num_node_p_switch", "2", "--num_gpu_p_node", "8", "--quiet", "--seed", "3", *extra])
    s = run(flags, output_dir=str(tmp_path / "o"))
    assert s["finished_jobs"] == 40 and s["queued_at_end"] == 0
    assert s["avg_jct"] > 0 and s["makespan"] > 0


def test_gpu_demands_mode(tmp_path):
    _, _, out = _sim(tmp_path, [(1, 0, 10), (2, 5, 10), (2, 6, 10)], "gpu-demands")
    rows = list(csv.reader(open(os.path.join(out, "job.csv"))))
    assert rows[0] == ["time", "1-GPU", "2-GPU", "4-GPU", "8-GPU", "12-GPU", "16-GPU", "24-GPU", "32-GPU"]
    demand = {int(r[0]): (int(r[1]), int(r[2])) for r in rows[1:]}
    assert demand[0] == (1, 0) and demand[5] == (1, 2) and demand[6] == (1, 4) and demand[10] == (0, 4)


def test_gittins_clean_wakes_every_delta_and_finishes(tmp_path):
    """--gittins_clean (ours; what the real-time controller executes): highest index first, a wake-up every
    gittins_delta instead of the reference's growing period. Two 4-GPU jobs on 4 GPUs: a long one that
    starts alone and a short one arriving later. By then the long job has attained more GPU-time than the
    short sample of the distribution, so its index has dropped to 0 and the fresh arrival preempts it."""
    rows, s, _ = _sim(tmp_path, [(4, 0, 1000), (4, 100, 50)], "gittins",
                      extra=("--gittins_desc", "--gittins_clean", "--gittins_delta", "40"))
    assert s["finished_jobs"] == 2
    assert int(rows[1]["end_time"]) == 150                       # ran at once for its 50 units
    assert int(rows[0]["preempt"]) == 1 and int(rows[0]["end_time"]) == 1050
    # the reference-faithful loop still completes the same trace
    rows2, s2, _ = _sim(tmp_path, [(4, 0, 1000), (4, 100, 50)], "gittins", extra=("--gittins_delta", "40"))
    assert s2["finished_jobs"] == 2
