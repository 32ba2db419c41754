"""CPU loopback trainer: the same parameter-server / worker protocol as
:class:`tiresias_b200.parallel.trainer.Trainer` (bucketed reduce-scatter of gradients onto PS shards, SGD
with momentum on the fp32 master shard, all-gather of the new parameters), on ``torch`` CPU tensors over a
``gloo`` group. It lets the worker pool, the controller, preemption and resume be tested without GPUs
 and doubles as the executable statement of the protocol the fused kernels implement.
"""
from __future__ import annotations

import time
from typing import Dict

import torch
import torch.distributed as dist
from torch import nn


class LoopbackTrainer:
    def __init__(self, model_name: str, rank: int, world: int, group=None, lr: float = 0.05,
                 momentum: float = 0.9, hidden: int = 64, step_sleep: float = 0.0, seed: int = 0):
        if model_name == "__explode__":                       # fault-injection hook for the pool's tests
            raise RuntimeError("injected failure while building the model")
        self.rank, self.world, self.group = rank, world, group
        self.lr, self.momentum = lr, momentum
        self.step_sleep = step_sleep
        torch.manual_seed(seed)
        width = {"resnet50": 48, "vgg16": 96, "transformer": 64}.get(model_name, hidden)
        self.model = nn.Sequential(nn.Linear(32, width), nn.ReLU(), nn.Linear(width, 8))
        n = sum(p.numel() for p in self.model.parameters())
        self.padded = (n + world - 1) // world * world
        self.shard = self.padded // world
        flat = torch.zeros(self.padded)
        flat[:n] = torch.cat([p.detach().reshape(-1) for p in self.model.parameters()])
        self.param_flat = flat
        self.master = flat[rank * self.shard:(rank + 1) * self.shard].clone()
        self.mom = torch.zeros(self.shard)
        self._steps = 0
        g = torch.Generator().manual_seed(1000 + rank)
        self.x = torch.randn(16, 32, generator=g)
        self.y = torch.randint(0, 8, (16,), generator=g)
        self._loss = float("nan")
        self.slowdown = lambda: 1.0               # co-location emulation hook (exec/worker.py)
        self._scatter_back()

    def _scatter_back(self) -> None:
        off = 0
        with torch.no_grad():
            for p in self.model.parameters():
                p.copy_(self.param_flat[off:off + p.numel()].view_as(p))
                off += p.numel()

    def step_resident(self) -> None:
        self.model.zero_grad()
        loss = nn.functional.cross_entropy(self.model(self.x), self.y)
        loss.backward()
        self._loss = float(loss.detach())
        grad = torch.zeros(self.padded)
        off = 0
        for p in self.model.parameters():
            grad[off:off + p.numel()] = p.grad.reshape(-1)
            off += p.numel()
        # push: reduce-scatter onto the PS shards
        mine = torch.zeros(self.shard)
        if self.world > 1:
            self._gloo_reduce_scatter(mine, grad)
        else:
            mine.copy_(grad)
        mine /= self.world
        # synchronous SGD on the shard
        self.mom.mul_(self.momentum).add_(mine)
        self.master.add_(self.mom, alpha=-self.lr)
        # pull: all-gather the new parameters
        if self.world > 1:
            parts = [torch.zeros(self.shard) for _ in range(self.world)]
            dist.all_gather(parts, self.master, group=self.group)
            self.param_flat = torch.cat(parts)
        else:
            self.param_flat = self.master.clone()
        self._scatter_back()
        self._steps += 1
        if self.step_sleep:
            time.sleep(self.step_sleep * self.slowdown())

    def _gloo_reduce_scatter(self, out: torch.Tensor, full: torch.Tensor) -> None:
        buf = full.clone()
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group)     # gloo has no reduce_scatter
        out.copy_(buf[self.rank * self.shard:(self.rank + 1) * self.shard])

    step = step_resident

    def loss(self) -> float:
        return self._loss

    def synchronize(self) -> None:
        pass

    def check_status(self) -> int:
        return 0

    def state_dict(self) -> Dict[str, torch.Tensor]:
        return {"master": self.master.clone(), "momentum": self.mom.clone(), "steps": torch.tensor(self._steps)}

    def load_state_dict(self, state: Dict[str, torch.Tensor]) -> None:
        self.master.copy_(state["master"])
        self.mom.copy_(state["momentum"])
        self._steps = int(state["steps"])
        if self.world > 1:
            parts = [torch.zeros(self.shard) for _ in range(self.world)]
            dist.all_gather(parts, self.master, group=self.group)
            self.param_flat = torch.cat(parts)
        else:
            self.param_flat = self.master.clone()
        self._scatter_back()

    def close(self) -> None:
        pass
