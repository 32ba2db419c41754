"""Execution backend: replays trace jobs on the GPUs of one 8xB200 NVSwitch box.

The reference prices a distributed job with an analytic formula (core/network/network_service.py) and
never runs anything; here a persistent pool of one worker process per GPU (``worker``) trains each
job's model on the GPU block the scheduler placed it on, a real-time ``controller`` makes the
scheduling decisions, and ``replay`` compares executed job completion times with the simulator's.
"""
