"""Synchronous-SGD job trainer for one NVSwitch domain: parameter-server/worker (fused kernels),
ring all-reduce (fused kernels) and the PyTorch+NCCL PS/worker baseline.

This is the execution backend behind every trace job. The reference only *prices* a job's
communication (core/network/network_service.py:3-39: ``(model_size/bandwidth + cross*latency) *
iterations * 2``) under the assumptions "Parameter Server / Worker", "Synchronize SGD"
(README.md:3-7) with one worker and one PS shard per GPU (infra/switch.py:97-107). Here the same
job runs for real:

``mode="ps"``    every bucket's push (reduce-scatter onto the PS shard), the SGD+momentum update
                of the fp32 master shard and the broadcast of the new bf16 parameters are one
                kernel (``tb200_ps_push_pull``), launched from a gradient hook on a side stream so
                it overlaps the rest of backward. Layers marked for *pull* skip the broadcast and
                gather their weights inside their first forward GEMM (``tb200_gemm_bf16_nt_pull``).
``mode="ring"``  the README's unimplemented all-reduce mode: in-kernel all-reduce per bucket followed by
                the stand-alone fused optimizer kernel on replicated state. ``allreduce_algo=1`` (default)
                is the direct two-shot algorithm — every rank reduces one slice from all peers and
                stores it to all peers, which is what an NVSwitch (every peer at full bandwidth) wants:
                634 GB/s bus bandwidth at 2 GPUs vs 529 for NCCL; ``allreduce_algo=0`` is the classic
                neighbour ring with 2(W-1) flag-synchronised steps.
``mode="nccl"``  the baseline named by BASELINE.json: bucketed ``reduce_scatter`` + torch SGD on the
                shard + ``all_gather`` through NCCL, same buckets, the STOCK network (nn.BatchNorm2d,
                cuDNN / cuBLAS everywhere), eager.
                ``nccl_same_compute=True`` is the like-for-like arm: the SAME fused network, gradient packing,
                optimizer kernel and CUDA-graph capture as ``ps``; only the communication differs (NCCL
                ``reduce_scatter`` -> ``tb200_sgd_momentum`` on the shard -> NCCL ``all_gather`` instead of the
                fused push/pull kernel). The gap between that arm and ``ps`` is what the in-kernel push/pull
                buys; the gap between the two NCCL arms is what the compute kernels buy.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch
from torch import nn

from ..models.zoo import ModelSpec, build as zoo_build
from ..ops import native
from ..ops.linear import TcLinear
from .symm import SymmetricWorld

_CAPTURE_LOCK = threading.Lock()

# largest per-GPU activation (rows) for which a row-sharded weight is pulled inside its GEMM instead of being
# broadcast by the push kernel (see Trainer.__init__); measured cut-off, profiles/comm_pull_r2.md
PULL_MAX_ROWS = 128

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)
_ALIGN = 64  # elements; keeps every parameter view 128-byte aligned inside the flat buffers


@dataclass
class Bucket:
    index: int
    offset: int                       # element offset in the flat buffers
    length: int                       # padded to a multiple of 8 * world
    params: List[Tuple[nn.Parameter, int]] = field(default_factory=list)   # (param, flat offset)
    pull_layer: Optional[TcLinear] = None
    pull_rows: int = 0                # rows of the pulled weight every PS rank owns (a multiple of 128)
    pull_sync: Optional[torch.Tensor] = None     # the pull GEMM's per-tile arrival counters / epochs
    pending: int = 0

    def shard(self, rank: int, world: int) -> Tuple[int, int]:
        n = self.length // world
        return self.offset + rank * n, n


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def plan_buckets(model: nn.Module, world: int, bucket_elems: int, pull_min_elems: int,
                 allow_pull: bool) -> Tuple[List[Bucket], int]:
    """Greedy bucketing in reverse registration order (~ the order gradients become ready).

    Large TcLinear weights get a bucket of their own so that they can be row-sharded over the PS ranks and
    pulled inside the GEMM. The shard of every rank is a whole number of 128-row GEMM tiles: the bucket is
    padded with zero rows up to ``world`` x that size (ResNet-50's 1000-row classifier occupies 1024 rows of the
    flat buffers at 2, 4 and 8 ranks), so no out_features has to divide anything.
    """
    pull_weights: Dict[int, TcLinear] = {}
    if allow_pull and world > 1:
        for m in model.modules():
            if isinstance(m, TcLinear) and m.weight.numel() >= pull_min_elems:
                pull_weights[id(m.weight)] = m
    buckets: List[Bucket] = []
    cur: Optional[Bucket] = None
    off = 0
    quantum = 8 * world

    def close() -> None:
        nonlocal cur, off
        if cur is not None:
            cur.length = _round_up(off - cur.offset, quantum)
            off = cur.offset + cur.length
            buckets.append(cur)
            cur = None

    for p in reversed([q for q in model.parameters() if q.requires_grad]):
        if id(p) in pull_weights:
            close()
            off = _round_up(off, max(_ALIGN, quantum))
            rows = _round_up(p.shape[0], 128 * world)
            b = Bucket(len(buckets), off, rows * p.shape[1], [(p, off)], pull_weights[id(p)], rows // world)
            off += b.length
            buckets.append(b)
            continue
        if cur is None:
            off = _round_up(off, max(_ALIGN, quantum))
            cur = Bucket(len(buckets), off, 0)
        off = _round_up(off, _ALIGN)
        cur.params.append((p, off))
        off += p.numel()
        if off - cur.offset >= bucket_elems:
            close()
    close()
    if len(buckets) >= native.MAX_BUCKETS - 1:
        raise ValueError(f"{len(buckets)} buckets exceed the flag pad; raise bucket_mb")
    return buckets, _round_up(off, max(_ALIGN, quantum))


class Trainer:
    """One worker (= one rank = one GPU) of a synchronous-SGD job."""

    def __init__(self, spec: ModelSpec, batch_size: int, mode: str = "ps",
                 world: Optional[SymmetricWorld] = None, lr: float = 0.1, momentum: float = 0.9,
                 weight_decay: float = 1e-4, bucket_mb: float = 8.0, use_graph: bool = True,
                 pull_gemm: bool = True, pull_min_mb: float = 4.0, comm_ctas: int = 32,
                 spin_budget: int = 50_000_000, seed: int = 0, device: Optional[torch.device] = None,
                 model: Optional[nn.Module] = None, allreduce_algo: Optional[int] = None, comm_group=None,
                 stem_pad8: bool = False, nvls: Optional[bool] = None, nccl_same_compute: bool = False):
        if mode not in ("ps", "ring", "nccl"):
            raise ValueError(f"unknown mode {mode!r}")
        self.spec, self.mode, self.batch_size = spec, mode, batch_size
        self.nccl_same_compute = bool(nccl_same_compute) and mode == "nccl"
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.symm = world if world is not None else SymmetricWorld.from_env(self.device)
        self.world, self.rank = self.symm.world, self.symm.rank
        self.use_graph = use_graph
        self.comm_ctas, self.spin_budget = comm_ctas, spin_budget
        # NVLS (in-switch reduction / multicast stores): None = the world's default (SymmetricWorld.
        # nvls_preferred: TB200_NVLS=1 and every GPU supports multicast objects); the buffers then come
        # from the VMM allocator
        self.nvls = (self.symm.nvls_preferred() if nvls is None else
                     (bool(nvls) and self.symm.nvls_available())) if mode != "nccl" else False
        self._allreduce_algo_arg = allreduce_algo
        self.allreduce_algo = 1 if allreduce_algo is None else allreduce_algo
        self.comm_group = comm_group          # NCCL group of this job's ranks (mode "nccl" only)
        self.stem_pad8 = stem_pad8
        self._lib = native.lib()

        # (seeding and model initialisation use the device's default generator, which a capturing tenant of this process
        # has registered with its graph: hold the capture lock - profiles/replay_r2b/README.md)
        with _CAPTURE_LOCK, torch.device(self.device):
            torch.manual_seed(seed)       # identical initial weights on every rank
            # the NCCL arm is the stock-PyTorch network (nn.BatchNorm2d / nn.Linear, cuDNN + cuBLAS)
            self.model = (model if model is not None else
                          zoo_build(spec, fused=(mode != "nccl" or self.nccl_same_compute))).to(self.device)
        self.model.train()
        # Pull layers (K2): the GEMM's copy-engine warps fetch every remote B tile once over NVLink while the tensor
        # cores start on the local shard (csrc/gemm_tcgen05.cu). The first version re-fetched a remote tile per
        # 128-row block of the activation and was only worth it for one row block; TB200_PULL_MAX_ROWS keeps that
        # cut-off adjustable (profiles/comm_pull_r2.md has both generations' numbers).
        rows_per_step = batch_size * (spec.sample_shape[0] if spec.kind == "tokens" else 1)
        max_rows = int(os.environ.get("TB200_PULL_MAX_ROWS", str(PULL_MAX_ROWS)))
        pull_min_mb = float(os.environ.get("TB200_PULL_MIN_MB", str(pull_min_mb)))
        self.buckets, self.total = plan_buckets(
            self.model, self.world, int(bucket_mb * (1 << 20) / 2), int(pull_min_mb * (1 << 20) / 2),
            allow_pull=(pull_gemm and mode == "ps" and rows_per_step <= max_rows))

        self.pack_grads = mode != "nccl" or self.nccl_same_compute
        self._grad_view: Dict[int, torch.Tensor] = {}
        self._alloc_state()
        self._flatten_model()
        self.set_hyper(lr, momentum, weight_decay)
        self._install_hooks()

        self.stream = torch.cuda.Stream(self.device)
        self.comm_stream = torch.cuda.Stream(self.device)
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._graph_failed = False
        self._eager_steps = 0                    # eager steps run by THIS object (warm-up before capture)
        self._steps = 0
        self.native_per_step = 0
        self.native_launches = 0
        self._pending_nccl: List[Tuple[Bucket, object]] = []
        self._phase_events = None
        self._alloc_io()
        # parameters were flattened / flags zeroed on the current stream: nobody (this rank's step stream, or a
        # peer's push kernel storing into our buffers) may run ahead of that
        torch.cuda.synchronize(self.device)
        if mode != "nccl":
            self.symm.barrier()
        self.stream.wait_stream(torch.cuda.current_stream(self.device))

    # ------------------------------------------------------------------------------------------
    # state
    # ------------------------------------------------------------------------------------------
    def _alloc_state(self) -> None:
        dev, n = self.device, self.total
        if self.mode == "nccl":
            self.param_flat = torch.zeros(n, dtype=torch.bfloat16, device=dev)
            self.grad_flat = torch.zeros(n, dtype=torch.bfloat16, device=dev)
        else:
            self.symm.alloc("param", n * 2, multicast=self.nvls)
            self.symm.alloc("grad", n * 2, multicast=self.nvls)
            self.symm.alloc("flags", native.FLAG_WORDS * 4)
            self.nvls = bool(self.symm.buffers["grad"].mc_ptr and self.symm.buffers["param"].mc_ptr)
            if self._allreduce_algo_arg is None and self.nvls:
                self.allreduce_algo = 2              # one-shot multimem.ld_reduce + multimem.st
            self.param_flat = self.symm.buffers["param"].view(torch.bfloat16)
            self.grad_flat = self.symm.buffers["grad"].view(torch.bfloat16)
            self._world_struct = self.symm.native_world("grad", "param", "flags")
            self._grad_ptrs = (C.c_void_p * native.MAX_WORLD)(
                *[self.symm.buffers["grad"].ptrs[r] for r in range(self.world)])
        replicated = self.mode == "ring"
        shard_total = n if replicated else n // self.world
        self.master = torch.zeros(shard_total, dtype=torch.float32, device=dev)
        self.mom = torch.zeros(shard_total, dtype=torch.float32, device=dev)
        # per-bucket offset of this rank's shard inside master/mom
        self._shard_base: List[int] = []
        acc = 0
        for b in self.buckets:
            self._shard_base.append(acc)
            acc += b.length if replicated else b.length // self.world
        self.hyper = torch.zeros(4, dtype=torch.float32, device=dev)
        self.epoch = torch.ones(1, dtype=torch.int32, device=dev)         # step number, starts at 1
        self.counters = torch.zeros(max(2, len(self.buckets)) + 2, dtype=torch.int32, device=dev)
        self.ar_seq = torch.zeros(2, dtype=torch.int32, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)

    def set_hyper(self, lr: float, momentum: float, weight_decay: float) -> None:
        self.lr, self.momentum_v, self.weight_decay = lr, momentum, weight_decay
        self.hyper.copy_(torch.tensor([lr, momentum, weight_decay, 1.0 / self.world],
                                      dtype=torch.float32))

    @staticmethod
    def _physical(p: torch.Tensor) -> torch.Tensor:
        """1-D copy of ``p`` in the element order it has in memory (NHWC for conv weights)."""
        if p.dim() == 4:
            return p.detach().permute(0, 2, 3, 1).reshape(-1)
        return p.detach().reshape(-1)

    def _repoint(self, p: nn.Parameter, off: int) -> None:
        n = p.numel()
        pv, gv = self.param_flat[off: off + n], self.grad_flat[off: off + n]
        if p.dim() == 4:
            o, i, h, w = p.shape
            p.data = pv.view(o, h, w, i).permute(0, 3, 1, 2)       # channels_last strides
            gview = gv.view(o, h, w, i).permute(0, 3, 1, 2)
        else:
            p.data = pv.view(p.shape)
            gview = gv.view(p.shape)
        self._grad_view[id(p)] = gview
        # nccl arm (DDP-style): gradients accumulate in place into bucket views. ps / ring: autograd
        # hands over a fresh tensor per parameter and one pack kernel per bucket gathers them.
        p.grad = gview if self.pack_grads is False else None

    def _flatten_model(self) -> None:
        """Move every parameter into the flat bf16 buffer (views); fp32 masters stay on the PS."""
        replicated = self.mode == "ring"
        if self.spec.kind == "image":
            self.model.to(memory_format=torch.channels_last)
        with torch.no_grad():
            for bi, b in enumerate(self.buckets):
                full32 = torch.zeros(b.length, dtype=torch.float32, device=self.device)
                for p, off in b.params:
                    full32[off - b.offset: off - b.offset + p.numel()] = self._physical(p).float()
                self.param_flat[b.offset: b.offset + b.length] = full32.to(torch.bfloat16)
                n = b.length if replicated else b.length // self.world
                lo = 0 if replicated else self.rank * n
                base = self._shard_base[bi]
                self.master[base: base + n] = full32[lo: lo + n]
                for p, off in b.params:
                    self._repoint(p, off)
            # everything that is not a trainable parameter (BN statistics...) follows to bf16
            for mod in self.model.modules():
                if getattr(mod, "keep_fp32_buffers", False):
                    continue
                for name, buf in list(mod._buffers.items()):
                    if buf is not None and buf.is_floating_point():
                        mod._buffers[name] = buf.to(torch.bfloat16)
        self.set_hyper(0.1, 0.9, 1e-4)
        for b in self.buckets:
            if b.pull_layer is not None:
                b.pull_sync = torch.zeros(native.PULL_SYNC_WORDS, dtype=torch.int32, device=self.device)
                b.pull_layer._pull_ctx = (self, b)

    def _alloc_io(self) -> None:
        dev, bs = self.device, self.batch_size
        # Optional: feed the ResNets 8 input channels (rgb + 5 zeros) so the stem convolution takes cuDNN's
        # 8-channel-aligned sm_100 path. Measured on B200 at batch 256 that path is SLOWER for the 7x7/2
        # stem (2.72 ms vs 1.47 ms for the 3-channel kernel), so it stays off; the kernel and the model
        # support are kept for convolution back-ends where it pays.
        self._pad8 = bool(self.stem_pad8) and self.spec.kind == "image" and hasattr(self.model, "_stem")
        # ResNets on our own convolution kernels take the stem input packed for the tensor cores (csrc/stem.cu):
        # the uint8 -> bf16 normalisation is fused into the pack
        self._packed_stem = (self.spec.kind == "image" and bool(getattr(self.model, "packed_stem", False))
                             and not self._pad8)
        if self.spec.kind == "image":
            self.x_dev = torch.zeros((bs, *self.spec.sample_shape), dtype=torch.uint8, device=dev)
            in_shape = (*self.spec.sample_shape[:-1], 8) if self._pad8 else self.spec.sample_shape
            if self._packed_stem:
                in_shape = native.stem_packed_shape(bs, *self.spec.sample_shape[:2])[1:]
            self.x_in = torch.zeros((bs, *in_shape), dtype=torch.bfloat16, device=dev)
        else:
            self.x_dev = torch.zeros((bs, *self.spec.sample_shape), dtype=torch.int64, device=dev)
            self.x_in = self.x_dev
        ydim = (bs,) if self.spec.kind == "image" else (bs, *self.spec.sample_shape)
        self.y_dev = torch.zeros(ydim, dtype=torch.int64, device=dev)
        self.loss_dev = torch.zeros(1, dtype=torch.float32, device=dev)
        # two pinned slots, so that step i+1 can be enqueued before step i's loss is read (fit())
        self.loss_host = torch.zeros(2, dtype=torch.float32).pin_memory()
        self._loss_done = [torch.cuda.Event(), torch.cuda.Event()]
        self._loss_slot = 0
        self.h2d_bytes = self.x_dev.numel() * self.x_dev.element_size() + self.y_dev.numel() * 8
        self.d2h_bytes = 4
        # input prefetch (prefetch() / fit()): the next batch lands in a staging pair on its own stream while
        # the current step computes; the step then pays a device-to-device copy (~10 us) instead of the PCIe
        # transfer (~0.7 ms for a 256 x 224 x 224 x 3 uint8 batch)
        self.x_stage = torch.empty_like(self.x_dev)
        self.y_stage = torch.empty_like(self.y_dev)
        self.copy_stream = torch.cuda.Stream(dev)
        self._stage_ready = torch.cuda.Event()
        self._stage_free = torch.cuda.Event()
        self._staged = False

    # ------------------------------------------------------------------------------------------
    # gradient hooks -> communication
    # ------------------------------------------------------------------------------------------
    def _install_hooks(self) -> None:
        for b in self.buckets:
            b.pending = len(b.params)
            for p, _ in b.params:
                p.register_post_accumulate_grad_hook(self._make_hook(b))

    def _make_hook(self, bucket: Bucket):
        def hook(_param: torch.Tensor) -> None:
            bucket.pending -= 1
            if bucket.pending == 0:
                bucket.pending = len(bucket.params)
                self._bucket_ready(bucket)
        return hook

    def _pack_bucket(self, b: Bucket) -> None:
        """Gather the bucket's per-parameter gradients into the flat symmetric buffer: one launch per 64
        tensors (tb200_pack_bf16). A gradient whose memory layout differs from the parameter's falls
        back to a strided torch copy."""
        items = []
        for p, off in b.params:
            g = p.grad
            if g is None:
                continue
            if g.dtype == torch.bfloat16 and g.stride() == p.stride():
                items.append((g.data_ptr(), self.grad_flat.data_ptr() + 2 * off, p.numel()))
            else:
                self._grad_view[id(p)].copy_(g)
            p.grad = None
        if items:
            arr = (native.PackItem * len(items))(*[native.PackItem(s, d, n) for s, d, n in items])
            native.check(self._lib.tb200_pack_bf16(arr, len(items), native.stream_ptr()))

    def _bucket_ready(self, b: Bucket) -> None:
        if self.pack_grads:
            self._pack_bucket(b)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.comm_stream.wait_event(ev)
        with torch.cuda.stream(self.comm_stream):
            if self.mode == "ps":
                self._launch_push(b)
            elif self.mode == "ring":
                self._launch_ring(b)
            else:
                self._launch_nccl(b)

    def _launch_push(self, b: Bucket) -> None:
        off, n = b.shard(self.rank, self.world)
        base = self._shard_base[b.index]
        native.check(self._lib.tb200_ps_push_pull(
            C.byref(self._world_struct), b.index, off, n, native.BF16,
            self.master.data_ptr() + 4 * base, self.mom.data_ptr() + 4 * base, self.hyper.data_ptr(),
            self.epoch.data_ptr(), self.counters.data_ptr() + 4 * b.index, self.status.data_ptr(),
            self.spin_budget, self.comm_ctas, 0 if b.pull_layer is not None else 1,
            native.stream_ptr(self.comm_stream)))

    def _launch_ring(self, b: Bucket) -> None:
        st = native.stream_ptr(self.comm_stream)
        if self.world > 1:
            ptrs = (C.c_void_p * native.MAX_WORLD)(
                *[self.symm.buffers["grad"].ptrs[r] + 2 * b.offset for r in range(self.world)])
            native.check(self._lib.tb200_allreduce(
                C.byref(self._world_struct), ptrs, b.length, native.BF16, 1.0, self.allreduce_algo,
                self.ar_seq.data_ptr(), self.status.data_ptr(), self.spin_budget, self.comm_ctas, st))
        base = self._shard_base[b.index]
        native.check(self._lib.tb200_sgd_momentum(
            self.master.data_ptr() + 4 * base, self.mom.data_ptr() + 4 * base,
            self.grad_flat.data_ptr() + 2 * b.offset, native.BF16,
            self.param_flat.data_ptr() + 2 * b.offset, native.BF16, b.length, self.hyper.data_ptr(), st))

    def _launch_nccl(self, b: Bucket) -> None:
        import torch.distributed as dist
        n = b.length // self.world
        g = self.grad_flat[b.offset: b.offset + b.length]
        if not hasattr(b, "_shard_grad"):
            b._shard_grad = torch.empty(n, dtype=torch.bfloat16, device=self.device)
            b._shard_param = torch.empty(n, dtype=torch.bfloat16, device=self.device)
        if dist.is_initialized() and self.world > 1:
            dist.reduce_scatter_tensor(b._shard_grad, g, op=dist.ReduceOp.SUM, group=self.comm_group)
        else:
            b._shard_grad.copy_(g[:n])
        base = self._shard_base[b.index]
        if self.nccl_same_compute:
            # same optimizer kernel as the ring / ps paths: only the collectives come from NCCL
            native.check(self._lib.tb200_sgd_momentum(
                self.master.data_ptr() + 4 * base, self.mom.data_ptr() + 4 * base, b._shard_grad.data_ptr(), native.BF16,
                b._shard_param.data_ptr(), native.BF16, n, self.hyper.data_ptr(), native.stream_ptr(self.comm_stream)))
            p = self.param_flat[b.offset: b.offset + b.length]
            if dist.is_initialized() and self.world > 1:
                dist.all_gather_into_tensor(p, b._shard_param, group=self.comm_group)
            else:
                p[:n].copy_(b._shard_param)
            return
        m, v = self.master[base: base + n], self.mom[base: base + n]
        # plain torch SGD(momentum, weight decay) on the fp32 master shard
        grad = b._shard_grad.float().mul_(1.0 / self.world).add_(m, alpha=self.weight_decay)
        v.mul_(self.momentum_v).add_(grad)
        m.add_(v, alpha=-self.lr)
        b._shard_param.copy_(m)
        p = self.param_flat[b.offset: b.offset + b.length]
        if dist.is_initialized() and self.world > 1:
            dist.all_gather_into_tensor(p, b._shard_param, group=self.comm_group)
        else:
            p[:n].copy_(b._shard_param)

    # ------------------------------------------------------------------------------------------
    # one step
    # ------------------------------------------------------------------------------------------
    def _preprocess(self) -> torch.Tensor:
        if self.spec.kind == "image":
            if self._packed_stem:
                native.stem7x7_pack(self.x_dev, IMAGENET_MEAN, IMAGENET_STD, out=self.x_in)
            elif self._pad8:
                native.u8_to_bf16_norm_pad8(self.x_dev, IMAGENET_MEAN, IMAGENET_STD, out=self.x_in)
            else:
                native.u8_to_bf16_norm(self.x_dev, IMAGENET_MEAN, IMAGENET_STD, out=self.x_in)
            return self.x_in.permute(0, 3, 1, 2)       # logical NCHW, physical NHWC
        return self.x_in

    def _mark(self, name: str) -> None:
        """Phase boundary for :meth:`profile_phases` (a CUDA event on the compute stream)."""
        if self._phase_events is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(torch.cuda.current_stream(self.device))
            self._phase_events.append((name, ev))

    def _device_step(self) -> None:
        self._mark("begin")
        x = self._preprocess()
        self._mark("preprocess")
        logits = self.model(x)
        loss = torch.nn.functional.cross_entropy(
            logits.float().reshape(-1, logits.shape[-1]), self.y_dev.reshape(-1))
        self.loss_dev.copy_(loss.detach().reshape(1))
        self._mark("forward")
        loss.backward()
        self._mark("backward")
        with torch.cuda.stream(self.comm_stream):
            if self.mode == "ps":
                native.check(self._lib.tb200_step_end(
                    C.byref(self._world_struct), len(self.buckets), self.epoch.data_ptr(),
                    self.status.data_ptr(), self.spin_budget, native.stream_ptr(self.comm_stream)))
        torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)
        if not self.pack_grads:
            self.grad_flat.zero_()         # in-place accumulation needs a clean slate; packing overwrites
        self._mark("comm_tail")

    def profile_phases(self, x_host: torch.Tensor, y_host: torch.Tensor, steps: int = 5) -> Dict[str, float]:
        """Per-phase device time of an *eager* step, in ms (CUDA events on the compute stream, averaged
        over ``steps``; with ``torch.distributed`` initialised the maximum over ranks): ``h2d`` input copy,
        ``preprocess``, ``forward`` (incl. loss), ``backward`` (the per-bucket push / all-reduce kernels
        overlap it on the communication stream), ``comm_tail`` (what communication is left exposed after
        backward: the wait for the last buckets + step_end), ``total``. The reference times only its whole
        loop with ``time.time()`` (core/scheduling/schedule.py:179,210-212)."""
        acc: Dict[str, float] = {}
        for _ in range(steps):
            self._phase_events = []
            with torch.cuda.stream(self.stream):
                self._mark("start")
                self.x_dev.copy_(x_host, non_blocking=True)
                self.y_dev.copy_(y_host, non_blocking=True)
                self._device_step()
                self._steps += 1
            self.synchronize()
            evs, self._phase_events = self._phase_events, None
            names = {"begin": "h2d"}
            for (_, a), (name, b) in zip(evs, evs[1:]):
                key = names.get(name, name)
                acc[key] = acc.get(key, 0.0) + a.elapsed_time(b)
            acc["total"] = acc.get("total", 0.0) + evs[0][1].elapsed_time(evs[-1][1])
        out = {k: v / steps for k, v in acc.items()}
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and self.world > 1 and self.symm.group is None:
            t = torch.tensor([out[k] for k in sorted(out)], device=self.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            out = dict(zip(sorted(out), t.tolist()))
        return out

    def _capture(self) -> None:
        # Jobs that share a GPU live in threads of one process (exec/worker.py, calib/interference.py): captures are
        # serialised, and thread-local capture mode keeps the other tenants' CUDA calls legal while one captures.
        with _CAPTURE_LOCK:
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph, stream=self.stream, capture_error_mode="thread_local"):
                self._device_step()

    def _advance(self) -> None:
        """Run one device-side step on ``self.stream`` (eager, or capture-then-replay)."""
        graphable = self.use_graph and (self.mode != "nccl" or self.nccl_same_compute) and not self._graph_failed
        # three eager steps IN THIS PROCESS first (a resumed job restores `_steps` but has run nothing here yet:
        # first calls bind contexts, set function attributes and size workspaces, none of which may happen
        # inside a capture - profiles/replay_r2/README.md)
        if graphable and self._graph is None and self._eager_steps >= 3:
            torch.cuda.synchronize(self.device)
            self.symm.barrier()
            before = native.launch_count()
            try:
                self._capture()
            except Exception:  # noqa: BLE001 - NCCL builds that cannot be captured: stay eager, say so
                if self.mode != "nccl":
                    raise
                self._graph, self._graph_failed = None, True
                torch.cuda.synchronize(self.device)
            self.native_per_step = native.launch_count() - before
        if graphable and self._graph is not None:
            self._graph.replay()
            self.native_launches += self.native_per_step      # kernel nodes replayed from the graph
        else:
            before = native.launch_count()
            self._device_step()
            self._eager_steps += 1
            self.native_per_step = native.launch_count() - before
            self.native_launches += self.native_per_step
        self._steps += 1

    def step(self, x_host: Optional[torch.Tensor] = None, y_host: Optional[torch.Tensor] = None) -> None:
        """Enqueue one training step on host tensors (pinned for asynchronous copies), or — called without
        arguments — on the batch a preceding :meth:`prefetch` staged. The loss of this step is read back
        with :meth:`loss` (device->host copy enqueued here)."""
        with torch.cuda.stream(self.stream):
            if x_host is None:
                if not self._staged:
                    raise RuntimeError("step() without a batch needs a preceding prefetch()")
                self.stream.wait_event(self._stage_ready)
                self.x_dev.copy_(self.x_stage, non_blocking=True)
                self.y_dev.copy_(self.y_stage, non_blocking=True)
                self._stage_free.record(self.stream)
                self._staged = False
            else:
                self.x_dev.copy_(x_host, non_blocking=True)
                self.y_dev.copy_(y_host, non_blocking=True)
            self._advance()
            self._loss_slot ^= 1
            self.loss_host[self._loss_slot:self._loss_slot + 1].copy_(self.loss_dev, non_blocking=True)
            self._loss_done[self._loss_slot].record(self.stream)

    def prefetch(self, x_host: torch.Tensor, y_host: torch.Tensor) -> None:
        """Start the host->device copy of the NEXT step's batch (pinned memory) on the copy stream; it
        overlaps whatever the compute stream is running. The next :meth:`step` without arguments uses it."""
        if self._staged:
            raise RuntimeError("a prefetched batch is already waiting; call step() first")
        cs = self.copy_stream
        cs.wait_event(self._stage_free)                 # the previous consumer has read the staging pair
        with torch.cuda.stream(cs):
            self.x_stage.copy_(x_host, non_blocking=True)
            self.y_stage.copy_(y_host, non_blocking=True)
            self._stage_ready.record(cs)
        self._staged = True

    def fit(self, batches):
        """Train on an iterable of pinned host batches ``(x, y)`` and yield every step's loss. The copy of
        batch i+1 is in flight while step i computes (double-buffered input), so a step costs its device
        time plus a device-to-device copy; every batch still crosses PCIe and every loss is read back —
        step i's loss is read (and yielded) while step i+1 is already running, so the host never leaves the
        GPU without queued work. Losses come out in step order, one per batch."""
        it = iter(batches)
        nxt = next(it, None)
        if nxt is None:
            return
        self.prefetch(*nxt)
        pending = None                       # slot holding the previous step's loss, not yet read
        while nxt is not None:
            self.step()
            slot = self._loss_slot
            nxt = next(it, None)
            if nxt is not None:
                self.prefetch(*nxt)
            if pending is not None:          # read step i-1's loss while step i runs: the GPU never idles
                yield self._read_loss(pending)
            pending = slot
        yield self._read_loss(pending)

    def _read_loss(self, slot: int) -> float:
        self._loss_done[slot].synchronize()
        return float(self.loss_host[slot])

    def step_resident(self) -> None:
        """One step on the batch already resident on the device (device-only timing)."""
        with torch.cuda.stream(self.stream):
            self._advance()

    def loss(self) -> float:
        """Blocks until the most recent step's loss has landed in pinned host memory."""
        self.stream.synchronize()
        return float(self.loss_host[self._loss_slot])

    def synchronize(self) -> None:
        self.stream.synchronize()
        self.comm_stream.synchronize()
        self.copy_stream.synchronize()

    def check_status(self) -> int:
        """Watchdog hook: non-zero means an in-kernel wait ran out of spin budget (peer lost)."""
        self.synchronize()
        return int(self.status.item())

    # ------------------------------------------------------------------------------------------
    # checkpoint / resume (preemption support for the executor)
    # ------------------------------------------------------------------------------------------
    def state_dict(self) -> Dict[str, torch.Tensor]:
        self.synchronize()
        out = {"master": self.master.detach().to("cpu", copy=True),
               "momentum": self.mom.detach().to("cpu", copy=True),
               "epoch": self.epoch.detach().to("cpu", copy=True),
               "steps": torch.tensor(self._steps)}
        for name, buf in self.model.named_buffers():
            out["buffer." + name] = buf.detach().to("cpu", copy=True)
        return out

    def load_state_dict(self, state: Dict[str, torch.Tensor]) -> None:
        """Restore the optimizer shard (fp32 masters + momentum), the step counter and the model buffers, then
        rebuild the FULL bf16 parameter buffer on every rank: in ``ps`` mode a rank only owns 1/world of the
        masters, so each rank writes its restored shard into every peer's parameter buffer over the symmetric
        mapping (what the push kernel's broadcast does at the end of a step). A barrier on both sides makes sure
        no rank starts a step on a mix of restored and freshly initialised weights, and that pull-GEMM layers
        (whose first wait, for epoch 0, passes at once) find their owners' shards in place."""
        self.synchronize()
        torch.cuda.synchronize(self.device)
        self.master.copy_(state["master"])
        self.mom.copy_(state["momentum"])
        self._steps = int(state["steps"])
        # the in-kernel flag protocol restarts at epoch 1 together with the fresh flag pads of this trainer (the
        # saved "epoch" is bookkeeping only): epochs only have to agree between the ranks of one incarnation
        bufs = dict(self.model.named_buffers())
        for k, v in state.items():
            if k.startswith("buffer."):
                bufs[k[len("buffer."):]].copy_(v)
        replicated = self.mode == "ring"
        with torch.no_grad():
            for bi, b in enumerate(self.buckets):
                n = b.length if replicated else b.length // self.world
                lo = 0 if replicated else self.rank * n
                base = self._shard_base[bi]
                shard = self.master[base: base + n].to(torch.bfloat16)
                self.param_flat[b.offset + lo: b.offset + lo + n] = shard
                if not replicated and self.world > 1:
                    if self.mode == "ps":
                        # peers' parameter buffers are mapped into this process: store the shard there too
                        from .symm import tensor_from_ptr
                        pbuf = self.symm.buffers["param"]
                        for r in range(self.world):
                            if r == self.rank:
                                continue
                            peer = tensor_from_ptr(pbuf.ptrs[r] + 2 * (b.offset + lo), 2 * n, torch.bfloat16, self.device)
                            peer.copy_(shard)
                    else:
                        import torch.distributed as dist
                        dist.all_gather_into_tensor(self.param_flat[b.offset: b.offset + b.length], shard,
                                                    group=self.comm_group)
        self._graph = None                       # captured before the restore: re-capture on the new state
        self._eager_steps = 0
        torch.cuda.synchronize(self.device)
        self.symm.barrier()
        self.stream.wait_stream(torch.cuda.current_stream(self.device))

    def close(self) -> None:
        self.synchronize()
        self._graph = None
        if self.device.type == "cuda":
            # per-stream kernel scratch of this job (a worker process outlives its jobs)
            for s in (self.stream, self.comm_stream, self.copy_stream):
                native.release_stream_caches(self.device, s)
        if self.mode != "nccl":
            self.symm.close()
