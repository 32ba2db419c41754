"""Execution backend without GPUs: buddy placement, the PS/worker protocol on CPU tensors over gloo
(world_size > 1), the worker pool + real-time controller, preemption and resume, and the simulated-vs-
executed JCT comparison."""
import os
import subprocess
import sys

import pytest
import torch

from tiresias_b200.exec.buddy import BuddyAllocator, all_blocks, normalize_gpus


def test_buddy_blocks_and_best_fit():
    assert normalize_gpus(12, 8) == 8 and normalize_gpus(3, 8) == 2 and normalize_gpus(1, 8) == 1
    assert len(all_blocks(8)) == 15
    b = BuddyAllocator(8)
    one = b.alloc(1, "a")
    assert one == (0, 1)
    assert b.alloc(1, "b") == (1, 1)            # packs next to the first single, keeping quads whole
    assert b.alloc(4, "c") == (4, 4)
    assert b.alloc(2, "d") == (2, 2)
    assert b.alloc(4, "e") is None and b.free_gpus() == 0 + 0
    b.free((4, 4))
    assert b.can_alloc(4) and not b.can_alloc(8)
    b.free((0, 1)); b.free((1, 1)); b.free((2, 2))
    assert b.alloc(8, "f") == (0, 8)


def _write_trace(path, jobs):
    with open(path, "w") as fh:
        fh.write("job_id,num_gpu,submit_time,iterations,model_name,duration,interval\n")
        for i, (g, s, it, m) in enumerate(jobs):
            fh.write(f"{i},{g},{s},{it},{m},10,0\n")
    return str(path)


@pytest.mark.timeout(300)
def test_replay_fifo_loopback_world2(tmp_path):
    from tiresias_b200.exec.replay import replay
    trace = _write_trace(tmp_path / "t.csv", [(2, 0, 30, "resnet50"), (1, 0, 20, "vgg16"), (1, 0, 20, "transformer"),
                                              (4, 1, 10, "resnet50")])
    res = replay(trace, gpus=2, schedule="fifo", comm="ps", out=str(tmp_path / "o"), loopback=True, step_sleep=0.01)
    recs = {r["job_id"]: r for r in res["records"]}
    assert all(r["status"] == "DONE" for r in recs.values())
    assert recs["3"]["num_gpu"] == 2                         # 4-GPU request clamps to the 2-GPU box
    assert recs["1"]["start"] >= recs["0"]["end"] - 0.05     # fifo: job 1 waits for the 2-GPU job ahead of it
    assert res["comparison"]["n"] == 4
    assert os.path.exists(tmp_path / "o" / "replay.json") and os.path.exists(tmp_path / "o" / "sim" / "jobs.csv")


@pytest.mark.timeout(300)
def test_replay_preempt_and_resume_loopback(tmp_path):
    from tiresias_b200.exec.replay import replay
    trace = _write_trace(tmp_path / "t.csv", [(2, 0, 200, "resnet50"), (1, 1, 20, "vgg16")])
    res = replay(trace, gpus=2, schedule="sjf", comm="ps", out=str(tmp_path / "o"), loopback=True, step_sleep=0.01,
                 check_every=5)
    recs = {r["job_id"]: r for r in res["records"]}
    assert recs["0"]["status"] == "DONE" and recs["1"]["status"] == "DONE"
    assert recs["0"]["preempt"] >= 1 and recs["0"]["attempts"] >= 2      # snapshot, requeue, resume
    assert recs["1"]["end"] < recs["0"]["end"]
    log = open(tmp_path / "o" / "controller.log").read()
    assert "(resume)" in log and "preempt job 0" in log


@pytest.mark.timeout(300)
def test_replay_dlas_gpu_loopback(tmp_path):
    """Executed discretised LAS: the long job is demoted to Q1 once its attained GPU-time passes the queue
    limit, so a later arrival (Q0) preempts it and — staying under the limit itself — runs to the end."""
    from tiresias_b200.exec.replay import replay
    trace = _write_trace(tmp_path / "t.csv", [(2, 0, 300, "resnet50"), (2, 1, 8, "vgg16")])
    res = replay(trace, gpus=2, schedule="dlas-gpu", comm="ps", out=str(tmp_path / "o"), loopback=True,
                 step_sleep=0.01, check_every=5, queue_limit_s=(0.5,))
    recs = {r["job_id"]: r for r in res["records"]}
    assert all(r["status"] == "DONE" for r in recs.values())
    assert recs["0"]["preempt"] >= 1 and recs["1"]["preempt"] == 0
    assert recs["1"]["end"] < recs["0"]["end"]
    sim = res["comparison"]
    assert sim["n"] == 2


@pytest.mark.timeout(300)
def test_replay_gittins_loopback(tmp_path):
    """Executed Gittins scheduling: a long job that has already received more GPU-time than the short jobs of
    the distribution ever need has a lower index than a fresh arrival, so it is preempted for it."""
    from tiresias_b200.exec.replay import replay
    trace = _write_trace(tmp_path / "t.csv", [(2, 0, 300, "resnet50"), (2, 1, 20, "vgg16"), (2, 1, 20, "vgg16")])
    res = replay(trace, gpus=2, schedule="gittins", comm="ps", out=str(tmp_path / "o"), loopback=True,
                 step_sleep=0.01, check_every=5, gittins_delta_s=0.2)
    recs = {r["job_id"]: r for r in res["records"]}
    assert all(r["status"] == "DONE" for r in recs.values())
    assert recs["0"]["preempt"] >= 1                         # the long job made way for a fresh arrival
    assert recs["2"]["end"] < recs["0"]["end"]               # which finished first
    assert res["comparison"]["n"] == 3 and os.path.exists(tmp_path / "o" / "sim" / "jobs.csv")


def test_loopback_trainer_matches_single_process_sgd():
    """world = 1: the PS protocol degenerates to plain SGD with momentum on the flat parameters."""
    from tiresias_b200.exec.loopback import LoopbackTrainer
    tr = LoopbackTrainer("resnet50", 0, 1, None, lr=0.05, momentum=0.9)
    ref = torch.nn.Sequential(torch.nn.Linear(32, 48), torch.nn.ReLU(), torch.nn.Linear(48, 8))
    ref.load_state_dict(tr.model.state_dict())
    opt = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    for _ in range(5):
        tr.step_resident()
        opt.zero_grad()
        torch.nn.functional.cross_entropy(ref(tr.x), tr.y).backward()
        opt.step()
    for a, b in zip(tr.model.parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=1e-6)
    state = tr.state_dict()
    tr2 = LoopbackTrainer("resnet50", 0, 1, None, seed=5)
    tr2.load_state_dict(state)
    for a, b in zip(tr2.model.parameters(), tr.model.parameters()):
        assert torch.equal(a, b)


@pytest.mark.timeout(300)
def test_failed_job_is_reported_and_the_pool_keeps_serving(tmp_path):
    """Fault injection: a job whose model cannot be built fails on its ranks; the controller marks it FAILED,
    frees its GPUs, and the jobs behind it still run (the reference aborts the whole simulation on any
    error, core/util.py:17-19)."""
    from tiresias_b200.exec import loopback
    from tiresias_b200.exec.replay import replay
    trace = _write_trace(tmp_path / "t.csv", [(2, 0, 10, "__explode__"), (1, 0, 10, "vgg16")])
    res = replay(trace, gpus=2, schedule="fifo", comm="ps", out=str(tmp_path / "o"), loopback=True, step_sleep=0.005)
    recs = {r["job_id"]: r for r in res["records"]}
    assert recs["0"]["status"] == "FAILED" and recs["1"]["status"] == "DONE"
    assert hasattr(loopback, "LoopbackTrainer")


def test_fd_exchange_between_ranks():
    """File descriptors (the handles of VMM / multicast allocations) travel between the ranks of a host
    as SCM_RIGHTS over unix sockets: three gloo ranks swap pipes and read each other's message."""
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=3",
                          "--master-addr", "127.0.0.1", "--master-port", str(port),
                          os.path.join(root, "tests", "fd_worker.py")], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0 and "FD OK" in res.stdout, (res.stdout + res.stderr)[-2000:]


@pytest.mark.timeout(300)
def test_replay_packed_single_gpu_jobs_loopback(tmp_path):
    """Executed co-location (reference: up to 3 tasks per device, infra/device.py:49-59): with one GPU slot left
    busy, a third single-GPU job joins a GPU that already runs one instead of waiting; the workers emulate the
    slowdown table on the CPU backend and the fifo-pack simulation of the same trace agrees with what ran."""
    from tiresias_b200.exec.replay import replay
    trace = _write_trace(tmp_path / "t.csv", [(1, 0, 150, "resnet50"), (1, 0, 150, "resnet50"), (1, 0.3, 40, "vgg16")])
    res = replay(trace, gpus=2, schedule="fifo", comm="ps", out=str(tmp_path / "o"), loopback=True, step_sleep=0.01,
                 pack=True, max_share=3, interference=[1.0, 1.5, 2.0])
    recs = {r["job_id"]: r for r in res["records"]}
    assert all(r["status"] == "DONE" for r in recs.values())
    assert recs["2"]["start"] < min(recs["0"]["end"], recs["1"]["end"])       # did not wait for a free GPU
    log = open(tmp_path / "o" / "controller.log").read()
    assert "shares the GPU with 1" in log
    # the host of the third job ran slower than the undisturbed one
    shared_with = "0" if recs["0"]["end"] > recs["1"]["end"] else "1"
    other = "1" if shared_with == "0" else "0"
    assert recs[shared_with]["end"] > recs[other]["end"] + 0.1
    assert res["pack"] and res["comparison"]["n"] == 3
    assert abs(res["comparison"]["avg_jct_error_pct"]) < 25.0


@pytest.mark.timeout(300)
def test_replay_gandiva_time_slicing_loopback(tmp_path):
    """Executed Gandiva time slicing (reference core/scheduling/algorithm.py:418-438): two jobs that both need the
    whole box take turns quantum by quantum; each is preempted at least once and both finish."""
    from tiresias_b200.exec.replay import replay
    trace = _write_trace(tmp_path / "t.csv", [(2, 0, 150, "resnet50"), (2, 0.2, 100, "vgg16")])
    res = replay(trace, gpus=2, schedule="gandiva", comm="ps", out=str(tmp_path / "o"), loopback=True, step_sleep=0.01,
                 check_every=5, quantum_s=0.5)
    recs = {r["job_id"]: r for r in res["records"]}
    assert all(r["status"] == "DONE" for r in recs.values())
    assert recs["0"]["preempt"] >= 1 and recs["1"]["preempt"] >= 1
    assert recs["1"]["start"] < recs["0"]["end"]                               # the second job did not wait for the first
    assert res["comparison"]["n"] == 2
