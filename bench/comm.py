#!/usr/bin/env python
"""Communication microbenchmark over real NVLink: the fused PS push/pull kernel and the in-kernel
all-reduce against their NCCL equivalents, with the achieved fraction of the NVLink roofline.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench/comm.py

Per bucket size (bf16 elements, the same on every rank):
  ours/ps     tb200_ps_push_pull  : every rank reduces its shard of all W gradient buffers over NVLink,
                                    applies SGD+momentum on its fp32 master shard and stores the new bf16
                                    parameters into all W parameter buffers (+ tb200_step_end)
  nccl/ps     reduce_scatter_tensor + torch SGD on the shard + all_gather_into_tensor
  ours/ring   tb200_allreduce algo 0 (ring) and algo 1 (direct two-shot)
  nccl/ar     all_reduce
  pull_gemm   tb200_gemm_bf16_nt_pull (K2): y = x . W^T with W row-sharded over the PS ranks — B tiles are TMA-loaded
              straight from the owners' memory inside the GEMM and written to the local copy on the way — against
              NCCL all_gather_into_tensor of the shards followed by a cuBLAS GEMM (``--op pull_gemm``)
Roofline: per GPU the push moves (W-1)/W * n * 2 bytes in and the pull the same amount out, each direction
at the measured 770 GB/s peer-copy bandwidth (B200_PROFILING.md); the HBM side (fp32 master + momentum,
16 bytes per shard element, plus the local bf16 traffic) runs at the measured copy bandwidth. Roofline
time = max(NVLink time, HBM time). CUDA-event timing, max over ranks, 5 warm-ups, 20 timed iterations.
"""
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tiresias_b200.ops import native  # noqa: E402
from tiresias_b200.parallel.symm import SymmetricWorld  # noqa: E402

NVLINK_GBS = 770.0


def hbm_gbs():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        return 6650.0


def timed(fn, dev, iters=20, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(dev)
    dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([s.elapsed_time(e) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def pull_gemm_cases(symm, ws, rank, world, dev, hbm):
    """K2 against its NCCL + cuBLAS composite on the FC shapes of the model table (core/models.py:9-11) and the
    transformer LM head. Roofline = max(2MNK at the measured cuBLAS peak, (W-1)/W of the weight over NVLink at
    770 GB/s, the weight once through HBM)."""
    lib = native.lib()
    try:
        tf = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"]
    except Exception:
        tf = 1400.0
    param = symm.buffers["param"].view(torch.bfloat16)
    epoch = torch.ones(1, dtype=torch.int32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    sync = torch.zeros(native.PULL_SYNC_WORDS, dtype=torch.int32, device=dev)
    # every owner's shard counts as written for all the steps this benchmark will run
    symm.post_flags("flags", native.flag_param_done(0, 0), native.MAX_WORLD, 1 << 30)
    torch.cuda.synchronize(dev)
    dist.barrier()
    rows = []
    for name, m, n, k in (("vgg fc6", 256, 4096, 25088), ("vgg fc7", 256, 4096, 4096), ("resnet50 fc", 256, 1000, 2048),
                          ("lm head", 4096, 32000, 1024), ("vgg fc6 b1024", 1024, 4096, 25088)):
        rpr = (n + 128 * world - 1) // (128 * world) * 128
        npad = rpr * world
        if npad * k > param.numel():
            continue
        w = param[: npad * k].view(npad, k)
        w.normal_(0, k ** -0.5)
        x = torch.randn(m, k, device=dev).to(torch.bfloat16)
        y = torch.empty(m, n, dtype=torch.bfloat16, device=dev)
        full = torch.empty(npad, k, dtype=torch.bfloat16, device=dev)
        shard = w[rank * rpr:(rank + 1) * rpr]
        torch.cuda.synchronize(dev)
        dist.barrier()

        def ours():
            native.check(lib.tb200_gemm_bf16_nt_pull(C.byref(ws), 0, 0, rpr, x.data_ptr(), y.data_ptr(), None, 0, m, n, k,
                                                     native.BF16, 0, epoch.data_ptr(), status.data_ptr(), 400_000_000,
                                                     sync.data_ptr(), native.stream_ptr()))
            epoch.add_(1)                              # a new step: the tiles have to be gathered again

        def nccl():
            dist.all_gather_into_tensor(full, shard.contiguous())
            torch.matmul(x, full[:n].t(), out=y)

        def local_ours():
            native.gemm_nt(x, w[:n], out=y)

        def local_cublas():
            torch.matmul(x, w[:n].t(), out=y)
        t_ours, t_nccl = timed(ours, dev), timed(nccl, dev)
        t_lo, t_lc = timed(local_ours, dev), timed(local_cublas, dev)
        link = (world - 1) / world * n * k * 2
        roof = max(2.0 * m * n * k / (tf * 1e9), link / (NVLINK_GBS * 1e6), (n * k * 2 + m * k * 2) / (hbm * 1e6))
        rows.append({"op": "pull_gemm", "shape": name, "M": m, "N": n, "K": k, "world": world, "rows_per_rank": rpr,
                     "ours_ms": round(t_ours, 4), "nccl_allgather_plus_cublas_ms": round(t_nccl, 4),
                     "speedup": round(t_nccl / t_ours, 3), "local_gemm_ours_ms": round(t_lo, 4),
                     "local_gemm_cublas_ms": round(t_lc, 4), "nvlink_rx_gbs": round(link / t_ours / 1e6, 1),
                     "roofline_ms": round(roof, 4), "frac_of_roofline": round(roof / t_ours, 3)})
    assert int(status.item()) == 0, "pull GEMM wait timed out"
    return rows


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--op", default="all", choices=["all", "ps", "pull_gemm"])
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    lib = native.lib()
    sizes = [1 << 20, 12 << 20, 48 << 20, 264 << 20]           # 2 MB .. 528 MB (VGG-16) of bf16 gradients
    nmax = max(sizes)
    symm = SymmetricWorld.from_env(dev)
    symm.alloc("grad", nmax * 2, multicast=True)       # NVLS view where the box has one
    symm.alloc("param", nmax * 2, multicast=True)
    symm.alloc("flags", native.FLAG_WORDS * 4)
    ws_mc = symm.native_world("grad", "param", "flags")
    has_mc = bool(ws_mc.grad_mc and ws_mc.param_mc)
    ws_p2p = symm.native_world("grad", "param", "flags")
    ws_p2p.grad_mc = None
    ws_p2p.param_mc = None
    ws = ws_p2p
    grad = symm.buffers["grad"].view(torch.bfloat16)
    param = symm.buffers["param"].view(torch.bfloat16)
    grad.normal_()
    master = torch.zeros(nmax // world, device=dev)
    mom = torch.zeros(nmax // world, device=dev)
    hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0 / world], device=dev)
    epoch = torch.ones(1, dtype=torch.int32, device=dev)
    counters = torch.zeros(8, dtype=torch.int32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    seq = torch.zeros(2, dtype=torch.int32, device=dev)
    ptrs = (C.c_void_p * native.MAX_WORLD)(*[symm.buffers["grad"].ptrs[r] for r in range(world)])
    st = native.stream_ptr()
    hbm = hbm_gbs()
    rows = []
    if args.op in ("all", "pull_gemm") and world > 1:
        rows += pull_gemm_cases(symm, ws_p2p, rank, world, dev, hbm)
    for n in (sizes if args.op in ("all", "ps") else []):
        shard = n // world
        for ctas, wsv, how in [(32, ws_p2p, "P2P"), (148, ws_p2p, "P2P")] + \
                ([(32, ws_mc, "NVLS"), (148, ws_mc, "NVLS")] if has_mc else []):
            def ours_ps():
                native.check(lib.tb200_ps_push_pull(C.byref(wsv), 0, rank * shard, shard, native.BF16,
                                                    master.data_ptr(), mom.data_ptr(), hyper.data_ptr(),
                                                    epoch.data_ptr(), counters.data_ptr(), status.data_ptr(),
                                                    400_000_000, ctas, 1, st))
                native.check(lib.tb200_step_end(C.byref(wsv), 1, epoch.data_ptr(), status.data_ptr(), 400_000_000, st))
            ms = timed(ours_ps, dev)
            # P2P: a GPU receives the other W-1 slices of its shard and sends its new shard W-1 times;
            # NVLS: the switch reduces / replicates, a GPU sends its W-1 foreign slices once and receives
            # the W-1 foreign shards once — the same bytes per direction, but no (W-1)-fold fan-in/out
            link_bytes = (world - 1) / world * n * 2                      # in (push) == out (pull), per direction
            hbm_bytes = shard * 16 + shard * 2 + shard * 2 + link_bytes  # master+mom r/w, local grad, local param, remote-served reads
            roof_ms = max(link_bytes / (NVLINK_GBS * 1e6), hbm_bytes / (hbm * 1e6)) if world > 1 else hbm_bytes / (hbm * 1e6)
            rows.append({"op": f"ps_push_pull+sgd (fused, {how})", "impl": "ours", "ctas": ctas, "elems": n, "world": world,
                         "ms": round(ms, 4), "nvlink_gbs_per_dir": round(link_bytes / ms / 1e6, 1),
                         "roofline_ms": round(roof_ms, 4), "frac_of_roofline": round(roof_ms / ms, 3)})
        g, p = grad[:n], param[:n]
        sg = torch.empty(shard, dtype=torch.bfloat16, device=dev)
        sp = torch.empty(shard, dtype=torch.bfloat16, device=dev)
        m, v = master[:shard], mom[:shard]

        def nccl_ps():
            if world > 1:
                dist.reduce_scatter_tensor(sg, g, op=dist.ReduceOp.SUM)
            else:
                sg.copy_(g[:shard])
            gr = sg.float().mul_(1.0 / world).add_(m, alpha=1e-4)
            v.mul_(0.9).add_(gr)
            m.add_(v, alpha=-0.1)
            sp.copy_(m)
            if world > 1:
                dist.all_gather_into_tensor(p, sp)
            else:
                p[:shard].copy_(sp)
        ms_n = timed(nccl_ps, dev)
        rows.append({"op": "reduce_scatter + torch SGD + all_gather", "impl": "nccl", "elems": n, "world": world,
                     "ms": round(ms_n, 4), "nvlink_gbs_per_dir": round((world - 1) / world * n * 2 / ms_n / 1e6, 1)})
        if world > 1:
            for algo, name in ((0, "ring"), (1, "direct")) + (((2, "NVLS one-shot"),) if has_mc else ()):
                wsa = ws_mc if algo == 2 else ws_p2p

                def ours_ar():
                    native.check(lib.tb200_allreduce(C.byref(wsa), ptrs, n, native.BF16, 1.0 / world, algo,
                                                     seq.data_ptr(), status.data_ptr(), 400_000_000, 64, st))
                ms_a = timed(ours_ar, dev)
                link = 2 * (world - 1) / world * n * 2
                rows.append({"op": f"allreduce ({name})", "impl": "ours", "elems": n, "world": world, "ms": round(ms_a, 4),
                             "busbw_gbs": round(link / ms_a / 1e6, 1),
                             "frac_of_roofline": round((link / 2 / (NVLINK_GBS * 1e6)) / ms_a, 3)})
            gg = g.clone()
            ms_b = timed(lambda: dist.all_reduce(gg), dev)
            rows.append({"op": "all_reduce", "impl": "nccl", "elems": n, "world": world, "ms": round(ms_b, 4),
                         "busbw_gbs": round(2 * (world - 1) / world * n * 2 / ms_b / 1e6, 1)})
        grad.normal_()
    torch.cuda.synchronize(dev)
    bad = int(status.item())
    if rank == 0:
        for r in rows:
            print(json.dumps(r), flush=True)
        print(json.dumps({"status_word": bad}))
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        tag = "" if args.op == "all" else "_" + args.op
        json.dump(rows, open(os.path.join(ROOT, "gpurun_out", f"comm_n{world}{tag}.json"), "w"), indent=1)
    dist.barrier()
    symm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
