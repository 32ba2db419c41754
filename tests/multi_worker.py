"""Worker body of tests/test_gpu_multi.py: one process per GPU under torchrun (NCCL bootstrap).

Checks, on real peer memory over NVLink:
  1. tb200_allreduce (ring and direct) against torch's sum of the per-rank inputs,
  2. the fused PS push/pull trainer against the PyTorch+NCCL PS/worker trainer (same buckets),
     eager and CUDA-graph replay, and that every rank ends with bit-identical parameters,
  3. the pull-GEMM path (weights gathered inside the first forward GEMM),
  4. ring all-reduce mode against the PS mode,
  5. checkpoint / resume at world > 1 (every rank restores only its PS shard; all ranks must continue from the
     full restored parameters) against an uninterrupted run, with and without pull-GEMM layers,
  6. the like-for-like NCCL arm (same fused compute, NCCL collectives), eager and captured in a CUDA graph,
  7. a randomised-delay stress of the flag / epoch protocol (the stand-in for compute-sanitizer's racecheck, which is
     closed on this GPU pool): every rank sleeps a different random time before each of 40 steps, so ranks run up to
     a whole step apart; reductions are in a fixed order, so ANY race shows up as parameters that are not bit-for-bit
     those of the undisturbed run.
Prints ``MULTI OK`` from rank 0 when everything held.
"""
import ctypes as C
import os
import sys

import torch
import torch.distributed as dist
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tiresias_b200.models.zoo import ModelSpec, synthetic_batch  # noqa: E402
from tiresias_b200.ops import native  # noqa: E402
from tiresias_b200.ops.linear import TcLinear  # noqa: E402
from tiresias_b200.parallel.symm import SymmetricWorld  # noqa: E402
from tiresias_b200.parallel.trainer import Trainer  # noqa: E402


class TinyNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv = nn.Conv2d(3, 16, 3, 2, 1, bias=False)
        self.bn = nn.BatchNorm2d(16)
        self.fc1 = TcLinear(16 * 8 * 8, 1024, act=native.ACT_RELU)
        self.fc2 = TcLinear(1024, 16)

    def forward(self, x):
        x = torch.relu(self.bn(self.conv(x)))
        x = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)
        return self.fc2(self.fc1(x))


TINY = ModelSpec("tiny", TinyNet, "image", (16, 16, 3), 16)


def check_allreduce(rank, world, dev):
    symm = SymmetricWorld.from_env(dev)
    n = 1 << 20
    symm.alloc("buf", n * 4, multicast=True)           # NVLS view where the box has one
    symm.alloc("flags", native.FLAG_WORDS * 4)
    ws = symm.native_world("buf", None, "flags")
    has_mc = bool(symm.buffers["buf"].mc_ptr)
    seq = torch.zeros(2, dtype=torch.int32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    lib = native.lib()
    for algo in (0, 1, 2):
        for dt, code in ((torch.float32, native.F32), (torch.bfloat16, native.BF16)):
            if algo == 2 and (not has_mc or dt != torch.bfloat16):
                continue                                   # NVLS all-reduce: bf16, needs the multicast view
            cnt = n if dt == torch.float32 else 2 * n
            view = symm.buffers["buf"].view(dt)
            g = torch.Generator(device=dev).manual_seed(17 * rank + algo)
            mine = torch.randn(cnt, device=dev, generator=g).to(dt)
            view.copy_(mine)
            gathered = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(gathered, mine)
            want = sum(t.float() for t in gathered) * 0.5
            torch.cuda.synchronize(dev)
            dist.barrier()
            ptrs = (C.c_void_p * native.MAX_WORLD)(*[symm.buffers["buf"].ptrs[r] for r in range(world)])
            native.check(lib.tb200_allreduce(C.byref(ws), ptrs, cnt, code, 0.5, algo, seq.data_ptr(),
                                             status.data_ptr(), 200_000_000, 32, native.stream_ptr()))
            torch.cuda.synchronize(dev)
            assert int(status.item()) == 0, f"allreduce algo {algo} timed out"
            got = view.float()
            tol = 1e-5 if dt == torch.float32 else 3e-2
            err = (got - want).abs().max().item()
            assert err <= tol * max(1.0, want.abs().max().item()), (algo, dt, err)
            dist.barrier()
    symm.close()
    return has_mc


def run_trainer(mode, rank, dev, steps=5, use_graph=False, pull=False, jitter=None, **kw):
    tr = Trainer(TINY, batch_size=32, mode=mode, use_graph=use_graph, device=dev, bucket_mb=0.05,
                 lr=0.05, pull_gemm=pull, pull_min_mb=0.25, **kw)
    losses = []
    rng = __import__("random").Random(jitter * 1000 + rank) if jitter is not None else None
    for i in range(steps):
        x, y = synthetic_batch(TINY, 32, seed=100 * rank + (i % 2))
        if rng is not None:
            __import__("time").sleep(rng.random() * 0.02)       # ranks drift up to a step apart
            tr.step(x, y)                                        # no loss read-back: nothing else keeps them in step
            continue
        tr.step(x, y)
        losses.append(tr.loss())
    assert tr.check_status() == 0, f"{mode}: in-kernel wait timed out"
    params = tr.param_flat.float().clone()
    n_pull = sum(1 for b in tr.buckets if b.pull_layer is not None)
    launches = tr.native_launches
    run_trainer.last_nvls = tr.nvls
    tr.close()
    return losses, params, n_pull, launches


def check_resume(rank, dev, world, pull):
    """Interrupt after 3 of 6 steps, restore into a trainer built from ANOTHER seed, compare with 6 straight steps."""
    kw = dict(batch_size=32, mode="ps", use_graph=False, device=dev, bucket_mb=0.05, lr=0.05, pull_gemm=pull,
              pull_min_mb=0.25)
    batches = [synthetic_batch(TINY, 32, seed=100 * rank + (i % 2)) for i in range(6)]
    tr = Trainer(TINY, **kw)
    for x, y in batches[:3]:
        tr.step(x, y)
    state = tr.state_dict()
    want = []
    for x, y in batches[3:]:
        tr.step(x, y)
        want.append(tr.loss())
    p_want = tr.param_flat.float().clone()
    tr.close()
    tr2 = Trainer(TINY, seed=4321, **kw)                       # different initial weights on purpose
    tr2.load_state_dict(state)
    got = []
    for x, y in batches[3:]:
        tr2.step(x, y)
        got.append(tr2.loss())
    assert tr2.check_status() == 0
    assert all(abs(a - b) < 5e-2 for a, b in zip(got, want)), (pull, got, want)
    if not pull:                                               # (a pulled weight's local copy lags by one forward)
        assert torch.allclose(tr2.param_flat.float(), p_want, rtol=2e-2, atol=2e-2)
        assert same_on_all_ranks(tr2.param_flat.float().clone(), world), "resume left ranks with different parameters"
    tr2.close()


def same_on_all_ranks(t, world):
    gathered = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(gathered, t)
    return all(torch.equal(gathered[0], g) for g in gathered[1:])


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

    has_mc = check_allreduce(rank, world, dev)

    l_ref, p_ref, _, _ = run_trainer("nccl", rank, dev)
    l_ps, p_ps, n_pull, launches = run_trainer("ps", rank, dev, nvls=False)       # peer loads / stores
    assert n_pull == 0 and launches > 0 and not run_trainer.last_nvls
    assert all(abs(a - b) < 5e-2 for a, b in zip(l_ps, l_ref)), (l_ps, l_ref)
    assert torch.allclose(p_ps, p_ref, rtol=2e-2, atol=2e-2), (p_ps - p_ref).abs().max().item()
    assert same_on_all_ranks(p_ps, world), "PS broadcast left ranks with different parameters"

    if has_mc:
        # the same job with the reduction done inside the NVSwitch and multicast parameter stores
        l_mc, p_mc, _, _ = run_trainer("ps", rank, dev, nvls=True)
        assert run_trainer.last_nvls, "NVLS available but the trainer did not use it"
        assert all(abs(a - b) < 5e-2 for a, b in zip(l_mc, l_ps)), (l_mc, l_ps)
        assert torch.allclose(p_mc, p_ps, rtol=2e-2, atol=2e-2), (p_mc - p_ps).abs().max().item()
        assert same_on_all_ranks(p_mc, world), "NVLS broadcast left ranks with different parameters"
        l_mg, p_mg, _, _ = run_trainer("ps", rank, dev, steps=8, use_graph=True, nvls=True)
        l_me, p_me, _, _ = run_trainer("ps", rank, dev, steps=8, use_graph=False, nvls=True)
        assert all(abs(a - b) < 5e-2 for a, b in zip(l_mg, l_me)), (l_mg, l_me)
        l_mr, p_mr, _, _ = run_trainer("ring", rank, dev, nvls=True)
        assert run_trainer.last_nvls
        assert torch.allclose(p_mr, p_ps, rtol=2e-2, atol=2e-2)
        assert same_on_all_ranks(p_mr, world), "NVLS all-reduce: ranks diverged"

    # pull-GEMM: fc1's weight is row-sharded over the PS ranks and gathered inside its GEMM.
    # The local copy of a pulled weight is refreshed by the GEMM's own TMA stores, so compare
    # after one more forward has run (run_trainer's last step did forward before its update).
    l_pl, p_pl, n_pull, _ = run_trainer("ps", rank, dev, pull=True)
    assert n_pull >= 1, "no layer qualified for the pull path"
    assert all(abs(a - b) < 5e-2 for a, b in zip(l_pl, l_ps)), (l_pl, l_ps)

    l_g, p_g, _, _ = run_trainer("ps", rank, dev, steps=8, use_graph=True)
    l_e, p_e, _, _ = run_trainer("ps", rank, dev, steps=8, use_graph=False)
    assert all(abs(a - b) < 5e-2 for a, b in zip(l_g, l_e)), (l_g, l_e)
    assert torch.allclose(p_g, p_e, rtol=2e-2, atol=2e-2)

    for algo in (0, 1):
        l_r, p_r, _, _ = run_trainer("ring", rank, dev, allreduce_algo=algo)
        assert all(abs(a - b) < 5e-2 for a, b in zip(l_r, l_ps)), (algo, l_r, l_ps)
        assert torch.allclose(p_r, p_ps, rtol=2e-2, atol=2e-2), algo
        assert same_on_all_ranks(p_r, world), f"ring algo {algo}: ranks diverged"

    # randomised-delay stress: bit-identical parameters with and without per-rank jitter, eager and graph, P2P
    # and (where present) NVLS; also with pull-GEMM layers, whose owners' shards are read while peers run ahead
    for kw in ({"nvls": False}, {"nvls": False, "use_graph": True}, {"nvls": False, "pull": True}) + \
            (({"nvls": True},) if has_mc else ()):
        _, p_calm, _, _ = run_trainer("ps", rank, dev, steps=40, **kw)
        _, p_jit, _, _ = run_trainer("ps", rank, dev, steps=40, jitter=7, **kw)
        assert torch.equal(p_calm, p_jit), ("timing changed the result", kw, (p_calm - p_jit).abs().max().item())
        assert same_on_all_ranks(p_jit, world) or kw.get("pull"), ("ranks diverged under jitter", kw)

    check_resume(rank, dev, world, pull=False)
    check_resume(rank, dev, world, pull=True)

    # like-for-like NCCL arm: same fused model / packing / optimizer kernel, NCCL collectives; eager and graph
    l_f, p_f, _, _ = run_trainer("nccl", rank, dev, nccl_same_compute=True)
    assert all(abs(a - b) < 5e-2 for a, b in zip(l_f, l_ps)), (l_f, l_ps)
    assert torch.allclose(p_f, p_ps, rtol=2e-2, atol=2e-2)
    l_fg, p_fg, _, _ = run_trainer("nccl", rank, dev, steps=8, use_graph=True, nccl_same_compute=True)
    assert all(abs(a - b) < 5e-2 for a, b in zip(l_fg, l_e)), (l_fg, l_e)

    dist.barrier()
    if rank == 0:
        print(f"MULTI OK (nvls={has_mc})", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
