"""The two entry points, as a user runs them (reference: run_sim.py:1741-1756, execute.py:44-53)."""
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def test_run_sim_subprocess_writes_reference_layout(tmp_path):
    trace = os.path.join(GOLD, "t12_yarn_fifo.trace.csv")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "run_sim.py"), "--trace_file", trace,
                          "--num_switch", "2", "--num_node_p_switch", "4", "--num_gpu_p_node", "4",
                          "--log_path", "cli_test", "--seed", "1", "--some_unknown_flag", "x"],
                         cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    runs = glob.glob(str(tmp_path / "log" / "cli_test" / "*"))
    assert len(runs) == 1 and len(os.path.basename(runs[0]).split("-")) == 7       # %Y-%m-%d-%H-%M-%S-%f
    names = sorted(os.listdir(runs[0]))
    assert names == ["cluster.csv", "cpu.csv", "gpu.csv", "job.csv", "jobs.csv", "memory.csv", "network.csv",
                     "output.log", "summary.json"]
    got = open(os.path.join(runs[0], "cluster.csv"), "rb").read()
    assert got == open(os.path.join(GOLD, "t12_yarn_fifo.cluster.csv"), "rb").read()
    assert "Total Time Taken in seconds" in open(os.path.join(runs[0], "output.log")).read()


def test_execute_in_process(tmp_path, monkeypatch):
    from tiresias_b200.cli import execute
    monkeypatch.chdir(tmp_path)
    trace = os.path.join(GOLD, "t12_yarn_fifo.trace.csv")
    execute.main(["--trace", trace, "--repeats", "2", "--config", "yarn:fifo:1", "--in-process",
                  "--num_switch", "2", "--num_node_p_switch", "4", "--quiet"])
    runs = glob.glob(str(tmp_path / "log" / "nodes_p_s4_job_*" / "yarn_fifo" / "*" / "cluster.csv"))
    assert len(runs) == 2


def test_demonstrate_trace_script(tmp_path):
    res = subprocess.run([sys.executable, os.path.join(ROOT, "demonstrate_trace.py"), "--no-plot"], cwd=tmp_path,
                         capture_output=True, text=True, timeout=120, env={**os.environ, "PYTHONPATH": ROOT})
    assert res.returncode == 0, res.stderr[-2000:]
    assert (tmp_path / "demonstrate_trace.csv").exists()
