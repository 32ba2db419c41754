"""Calibration of the simulator's cost model from measurements on the 8xB200 box: per-model iteration and
start-up times at 1/2/4/8 GPUs (replacing core/network/network_service.py's analytic cost) and the
co-location slowdown of 2-3 jobs sharing one GPU (replacing infra/interference.py's FACTOR = 0.2)."""
