"""Hook for a learned scheduling agent (reference stub: model/env.py:1-7, O4; ``Scheduler.agent``,
core/scheduling/schedule.py:25-27).

The reference leaves ``step`` empty. Here the environment wraps a :class:`TickScheduler`: an action
is the index of the queued job to try next; the observation is the cluster row of that tick; the
reward is minus the number of queued jobs (a proxy for pending time).
"""
from __future__ import annotations


class Environment:
    def __init__(self, scheduler=None):
        self.scheduler = scheduler

    def observe(self):
        if self.scheduler is None:
            return None
        return self.scheduler.construct_info()

    def step(self, action=None):
        """Advance the wrapped simulation by one tick. Returns (observation, reward, done)."""
        if self.scheduler is None:
            return None
        done = not self.scheduler.tick(action)
        info = self.scheduler.last_info
        return info, -float(info.num_queuing_jobs) if info is not None else 0.0, done
