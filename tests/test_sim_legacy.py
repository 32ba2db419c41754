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


# ---- decision oracles for the rest of the Tiresias family (each schedule worked out by hand from the rules of
# ---- the reference's dead loops; the comments carry the derivation) ----------------------------------------------

def test_lpjf_serves_longest_waiting_and_stops_at_first_failure(tmp_path):
    # run_sim.py:1474-1531. j0 2G@0 x100, j1 4G@10 x50, j2 1G@20 x10 on 4 GPUs. Arrivals always queue; the queue is
    # sorted by waiting time (longest first) and served until the first job that does not fit.
    # t=0  [j0] -> runs 0-100.   t=10 [j1]: 4 GPUs wanted, 2 free -> stop.
    # t=20 [j1 (waited 10), j2 (0)]: j1 fails -> stop; j2 is NOT tried although 2 GPUs are free
    #      (fjf, sorted by num_gpu, would start it at 20).
    # t=100 j0 ends: j1 fits -> 100-150 (removing it while iterating skips j2, which would not fit anyway).
    # t=150 j2 -> 150-160.
    rows, s, _ = _sim(tmp_path, [(2, 0, 100), (4, 10, 50), (1, 20, 10)], "lpjf")
    assert _ints(rows[0], "start_time", "end_time") == (0, 100)
    assert _ints(rows[1], "start_time", "end_time", "pending_time") == (100, 150, 90)
    assert _ints(rows[2], "start_time", "end_time", "JCT", "pending_time") == (150, 160, 140, 130)
    assert s["makespan"] == 160


def _dist(tmp_path, samples):
    path = tmp_path / "dist.csv"
    path.write_text("duration\n" + "".join(f"{v}\n" for v in samples))
    return str(path)


def test_gittins_faithful_ascending_sort_vs_descending(tmp_path):
    # Distribution {10, 20, 100, 200}, delta 50. Index table (run_sim.py:1649-1680), a = sample - 1:
    #   a=9  : idx 0, idx_delta 2 -> p = 2/4,   e = (10+20+50*2)/4 = 32.5 -> 15384.6154
    #   a=19 : idx 1, idx_delta 2 -> p = .33333, e = (20+50*2)/3   = 40   ->  8333.25
    #   a=99 : idx 2, idx_delta 3 -> p = 1/2,   e = (100+50)/2     = 75   ->  6666.6667
    #   a=199: past the last sample                                        ->     0
    # j0 4G@0 x300, j1 4G@100 x30 on 4 GPUs. The reference's loop (run_sim.py:948-1202) sorts ASCENDING (:1082):
    # at t=100 j0 has attained 400 GPU-s (rank 0) and the fresh j1 ranks 15384.6, so j0 sorts first and keeps the
    # GPUs; j1 only starts when j0 ends. Wake-ups: next_unit = 50, then += now each pass: 50, 100, 200, (400).
    dist = _dist(tmp_path, [10, 20, 100, 200])
    from tiresias_b200.sim.legacy import gittins as gi
    table = gi.parse_job_dist(dist, 50)
    assert table["gittins"][:4] == [15384.6154, 8333.25, 6666.6667, 0.0]
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    rows, s, out = _sim(tmp_path / "a", [(4, 0, 300), (4, 100, 30)], "gittins",
                        extra=("--job_dist", dist, "--gittins_delta", "50"))
    assert _ints(rows[0], "start_time", "end_time", "preempt") == (0, 300, 0)
    assert _ints(rows[1], "start_time", "end_time", "pending_time") == (300, 330, 200)
    times = [int(r["time"]) for r in csv.DictReader(open(os.path.join(out, "cluster.csv")))]
    assert times == [0, 50, 100, 200, 300, 330]
    # --gittins_desc (highest index first, the order of the reference's dlas-gpu-gittins variant, :809): the
    # fresh job preempts at once.
    rows, s, _ = _sim(tmp_path / "b", [(4, 0, 300), (4, 100, 30)], "gittins",
                      extra=("--job_dist", dist, "--gittins_delta", "50", "--gittins_desc"))
    assert _ints(rows[1], "start_time", "end_time") == (100, 130)
    assert _ints(rows[0], "end_time", "preempt", "pending_time") == (330, 1, 30)


def test_multi_dlas_reserves_gpus_per_class(tmp_path):
    # run_sim.py:432-661. 10 GPUs; j0 1G@0 x100, j1 4G@0 x100, j2 4G@10 x50. At t=0 the two new classes split the
    # cluster by demand (1 + 4 = 5 GPUs wanted): 1-GPU class floor(1/5*10) -> min(demand 1, 2) = 1 GPU, 4-GPU class
    # floor(8/4)*4 -> min(demand 4, 8) = 4 GPUs; the 5 spare GPUs stay unreserved because no class has unmet
    # demand. j2 arrives into a known class (no re-balance before t=600): its class owns 4 GPUs, all used by j1,
    # so it waits for j1 (100) although 5 GPUs are idle. Plain dlas-gpu, which counts the whole cluster, starts
    # it on arrival.
    jobs = [(1, 0, 100), (4, 0, 100), (4, 10, 50)]
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    rows, s, _ = _sim(tmp_path / "a", jobs, "multi-dlas-gpu", gpus=10)
    assert _ints(rows[2], "start_time", "end_time", "pending_time") == (100, 150, 90)
    assert s["makespan"] == 150
    rows, s, _ = _sim(tmp_path / "b", jobs, "dlas-gpu", gpus=10)
    assert _ints(rows[2], "start_time", "end_time") == (10, 60)


def test_dlas_gpu_pack_shares_gpus_by_model_memory(tmp_path):
    # run_sim.py:1204-1472 + infra/cluster.py:117-148: a job takes model.mem_util of each of its num_gpu GPUs. The
    # reference's live table is all ones (core/models.py:22), so packing never engages and the schedule equals
    # dlas-gpu; with its commented-out measured table (core/models.py:21; resnet50 = 0.50, --pack_mem_util table)
    # two resnet50 jobs share both GPUs of a 2-GPU cluster and the second starts on arrival.
    jobs = [(2, 0, 100), (2, 10, 50)]
    for name in "abc":
        (tmp_path / name).mkdir()
    rows, s, _ = _sim(tmp_path / "a", jobs, "dlas-gpu-pack", gpus=2, extra=("--pack_mem_util", "table"))
    assert _ints(rows[1], "start_time", "end_time") == (10, 60) and s["makespan"] == 100
    rows, s, _ = _sim(tmp_path / "b", jobs, "dlas-gpu-pack", gpus=2)
    assert _ints(rows[1], "start_time", "end_time") == (100, 150) and s["makespan"] == 150
    rows_plain, _, _ = _sim(tmp_path / "c", jobs, "dlas-gpu", gpus=2)
    assert _ints(rows_plain[1], "start_time", "end_time") == (100, 150)


def test_gandiva_legacy_time_slices_an_oversubscribed_node_set(tmp_path):
    # run_sim.py:101-158 + infra/cluster.py:386-489. One 4-GPU node, two 4-GPU jobs of 100 s at t=0: the node set
    # of the 4-GPU class has capacity 1, the second job over-subscribes it. The loop steps 10 s; only the head job
    # advances; heads rotate when the clock hits a multiple of 60:
    #   j0 advances at t=10..60 (60 s done), rotate at 60; j1 advances t=70..120 (60 s), rotate at 120;
    #   j0 advances t=130..160 -> 100 s, done at 160; j1 advances t=170..200 -> done at 200.
    # (fifo would finish j0 at 100.)
    rows, s, _ = _sim(tmp_path, [(4, 0, 100), (4, 0, 100)], "gandiva-legacy", scheme="gandiva", gpus=4)
    assert _ints(rows[0], "start_time", "end_time", "JCT") == (0, 160, 160)
    assert _ints(rows[1], "start_time", "end_time", "JCT") == (0, 200, 200)
    assert s["makespan"] == 200


def test_lp_placement_is_optimal_against_brute_force():
    # core/lp.py:2-44: minimise the maximum per-node network load; binary p[j, n]; GPU / CPU capacity; every
    # component placed once. 2 PS shards + 2 workers on 3 nodes with uneven existing load: 81 assignments.
    import itertools
    import numpy as np
    from tiresias_b200.sim.legacy import lp
    info = {"t": [30.0, 12.5], "num_ps": 2, "num_w": 2, "m_size": 42.5, "ps_c": lp.PS_CPU, "w_c": lp.WORKER_CPU}
    cl = {"fg": [1, 1, 2], "fc": [16, 6, 16], "ct": [20.0, 0.0, 55.0], "num_n": 3}
    x = lp._solve_milp(info, cl)
    assert x is not None and x.shape == (4, 3) and (x.sum(1) == 1).all()
    t = [info["t"][0] * 2, info["t"][1] * 2, info["m_size"], info["m_size"]]
    g, c = [0, 0, 1, 1], [lp.PS_CPU, lp.PS_CPU, lp.WORKER_CPU, lp.WORKER_CPU]

    def objective(assign):
        load, fg, fc = list(cl["ct"]), list(cl["fg"]), list(cl["fc"])
        for j, n in enumerate(assign):
            load[n] += t[j]; fg[n] -= g[j]; fc[n] -= c[j]
        return max(load) if min(fg) >= 0 and min(fc) >= 0 else None

    feasible = [v for v in (objective(a) for a in itertools.product(range(3), repeat=4)) if v is not None]
    got = objective([int(np.argmax(x[j])) for j in range(4)])
    assert got is not None and abs(got - min(feasible)) < 1e-6, (got, min(feasible))
    # capacity is honoured: node 1 has CPUs for either one worker or one PS shard, never both workers
    assert x[2:, 1].sum() <= 1


def test_cross_node_job_network_load_follows_the_switch_formula(tmp_path):
    # infra/switch.py:97-107 / :122-133: every GPU hosts one worker and one PS shard; per node
    #   traffic = model_size * local_gpus + sum_{local PS i} ps[i] * (need_gpu - local_gpus) - ps[i] * local_gpus
    # A 4-GPU resnet50 job on 2 nodes x 2 GPUs under yarn placement: 2 workers + shards (0,1) / (2,3) per node.
    from tiresias_b200.sim.legacy import engine
    trace = tmp_path / "t.csv"
    trace.write_text(HEADER + "0,4,0,100,resnet50,100,30\n")
    flags = define_simulator_flags()
    flags.parse(["--trace_file", str(trace), "--schedule", "fjf", "--scheme", "yarn", "--num_switch", "1",
                 "--num_node_p_switch", "2", "--num_gpu_p_node", "2", "--quiet"])
    (tmp_path / "o").mkdir()
    sim = engine.build(flags, str(tmp_path / "o"))
    job = sim.jobs.job_list[0]
    assert sim.cluster.try_get_job_res(job)
    nodes = [nd for pl in job.placements for nd in pl["nodes"]]
    assert [nd["num_gpu"] for nd in nodes] == [2, 2]
    size, ps = job.model["total_size"], job.ps_network
    # resnet50's tensors dealt round-robin over 4 shards (core/models.py:8-20, Job docstring core/jobs/job.py:61-73)
    shards = [0.0] * 4
    for i, tsize in enumerate(job.model["tensors"]):
        shards[i % 4] += tsize
    assert ps == [round(v, 1) for v in shards] and abs(sum(ps) - size) < 0.5
    want = []
    for first in (0, 2):
        traffic = size * 2
        for i in (first, first + 1):
            traffic += ps[i] * (4 - 2) - ps[i] * 2
        want.append(round(traffic, 1))
    assert [nd["network"] for nd in nodes] == pytest.approx(want, abs=0.2)
    assert want[0] == pytest.approx(size * 2, abs=0.2)      # shard terms cancel when half the workers are local


def test_fifo_pack_shares_a_gpu_at_the_measured_slowdown(tmp_path):
    # The executor's packing mode, simulated (sim/legacy/schedulers.py::fifo_shared). 2 GPUs, slowdown table
    # [1, 1.5, 2]: j0, j1 (1 GPU, 100 s) take a GPU each at t=0; j2 (1 GPU, 30 s) arrives at 10 with no GPU free and
    # joins the shared GPU with the fewest tenants (GPU 0, j0). Both then advance at 1/1.5: j2 needs 45 s -> ends 55;
    # j0 has done 10 + 45/1.5 = 40 by then and finishes its other 60 alone at 115. j1, undisturbed, ends at 100.
    jobs = [(1, 0, 100), (1, 0, 100), (1, 10, 30)]
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    rows, s, _ = _sim(tmp_path / "a", jobs, "fifo-pack", gpus=2, extra=("--pack_slowdown", "1,1.5,2"))
    assert _ints(rows[2], "start_time", "end_time") == (10, 55)
    assert _ints(rows[0], "end_time") == (115,) and _ints(rows[1], "end_time") == (100,)
    # one tenant per GPU: the same loop is plain fifo and the late job waits for a GPU
    rows, s, _ = _sim(tmp_path / "b", jobs, "fifo-pack", gpus=2, extra=("--pack_max_share", "1"))
    assert _ints(rows[2], "start_time", "end_time") == (100, 130)


def test_gandiva_slice_round_robins_while_jobs_wait(tmp_path):
    # The executor's gandiva policy, simulated: 2 GPUs, quantum 30. j0 (2 GPUs, 100 s) starts at 0; j1 (2 GPUs, 40 s)
    # arrives at 5 and waits. t=30: j0 has held the GPUs for a quantum -> preempted (30 done), j1 runs; t=60: j1
    # (30 done) yields to j0; t=90: j0 (60 done) yields to j1, which needs 10 more -> done at 100; j0 then runs its
    # last 40 s undisturbed (nothing waits, no slicing) -> 140.
    rows, s, _ = _sim(tmp_path, [(2, 0, 100), (2, 5, 40)], "gandiva-slice", gpus=2, extra=("--slice_quantum", "30"))
    assert _ints(rows[1], "start_time", "end_time", "preempt") == (30, 100, 1)
    assert _ints(rows[0], "start_time", "end_time", "preempt") == (0, 140, 2)
