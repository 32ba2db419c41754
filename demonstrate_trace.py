#!/usr/bin/env python
"""Draws 200 samples from the synthetic job generator, checks every field is populated and plots
the four empirical CDFs (reference: demonstrate_trace.py:1-66, C15). Without matplotlib, or with
``--no-plot``, the CDF points are written to ``demonstrate_trace.csv`` instead of a window."""
import sys

import numpy as np

from tiresias_b200.sim.trace import JobGenerator


def main(number: int = 200, plot: bool = True, debug: bool = False) -> dict:
    jg = JobGenerator()
    trace = jg.generate_trace(number)
    for field in ("model", "itter", "arrival", "duration", "num_gpu", "interval"):
        assert any(trace[field]), field
    if debug:
        print(trace)
    panels = [("Model Sizes", "model"), ("Itterations", "itter"), ("Job Arrival", "arrival"),
              ("Duration (Hours)", "duration")]
    curves = {title: (np.sort(trace[key]), jg.cdf(trace[key])) for title, key in panels}
    try:
        if not plot:
            raise ImportError
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        plt.figure()
        plt.suptitle("CDF ({} generated samples)".format(number))
        for pos, (title, (xs, ys)) in zip((221, 222, 223, 224), curves.items()):
            ax = plt.subplot(pos)
            ax.plot(xs, ys)
            ax.set_title(title)
        plt.savefig("demonstrate_trace.png")
    except ImportError:
        with open("demonstrate_trace.csv", "w") as fh:
            fh.write("panel,value,cdf\n")
            for title, (xs, ys) in curves.items():
                for x, y in zip(xs, ys):
                    fh.write(f"{title},{x},{y}\n")
    return trace


if __name__ == "__main__":
    main(plot="--no-plot" not in sys.argv, debug="--debug" in sys.argv)
