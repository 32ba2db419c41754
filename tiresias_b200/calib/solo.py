"""One single-GPU training process for the co-location measurement (see calib.interference)."""
from __future__ import annotations

import argparse
import json
import os
import time

import torch


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", required=True)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--iters", type=int, default=60)
    ap.add_argument("--mode", default="ps")
    ap.add_argument("--gpu", type=int, default=0)
    ap.add_argument("--go-file", required=True, help="start the timed region once this file exists")
    ap.add_argument("--out", required=True)
    a = ap.parse_args()
    from ..models import zoo
    from ..parallel.trainer import Trainer
    torch.cuda.set_device(a.gpu)
    dev = torch.device("cuda", a.gpu)
    spec = zoo.resolve(a.model)
    tr = Trainer(spec, a.batch, mode=a.mode, use_graph=True, device=dev)
    x, y = zoo.synthetic_batch(spec, a.batch)
    for _ in range(6):                      # allocator growth, graph capture
        tr.step(x, y)
    tr.loss()
    open(a.out + ".ready", "w").write("1")
    while not os.path.exists(a.go_file):
        time.sleep(0.005)
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    with torch.cuda.stream(tr.stream):
        start.record()
    for _ in range(a.iters):
        tr.step_resident()
    with torch.cuda.stream(tr.stream):
        end.record()
    tr.synchronize()
    wall = time.time() - t0
    res = {"model": a.model, "iters": a.iters, "iter_ms": start.elapsed_time(end) / a.iters,
           "wall_iter_ms": wall / a.iters * 1e3, "peak_mem_bytes": int(torch.cuda.max_memory_allocated(dev))}
    tr.close()
    with open(a.out, "w") as fh:
        json.dump(res, fh)


if __name__ == "__main__":
    main()
