"""bench.py's `sim` block: the unmodified reference (baseline/_ref, its own run_sim.py) against our tick engine."""
import pytest

from tiresias_b200.utils import refsim


def test_reference_and_ours_write_the_same_cluster_csv():
    why = refsim.reference_available()
    if why:
        pytest.skip(why)
    case = {"name": "t", "scheme": "yarn", "schedule": "fifo", "num_queue": 1, "num_jobs": 16, "seed": 3, "cluster": (2, 4, 8)}
    out = refsim.compare([case])
    row = out["cases"][0]
    assert "reference_error" not in row and "ours_error" not in row, row
    assert row["makespan_equal"] and row["cluster_csv_identical"], row
    assert row["reference_wall_s"] > 0 and row["ours_wall_s"] > 0


def test_compare_reports_unavailable_reference_without_failing(monkeypatch):
    monkeypatch.setattr(refsim, "reference_available", lambda: "forced by the test")
    case = {"name": "t", "scheme": "yarn", "schedule": "fifo", "num_queue": 1, "num_jobs": 8, "seed": 3, "cluster": (1, 2, 8)}
    out = refsim.compare([case])
    assert out["reference_unavailable"] == "forced by the test"
    assert "ours_makespan" in out["cases"][0] and "reference_wall_s" not in out["cases"][0]
