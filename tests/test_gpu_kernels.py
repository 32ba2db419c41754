"""Numerics of every hand-written kernel against a plain PyTorch fp32 reference of the same op
(SURVEY.md §4 item 5)."""
import ctypes as C
import os

import pytest
import torch

from tiresias_b200.ops import native

pytestmark = pytest.mark.gpu


def _bf16(t):
    return t.to(torch.bfloat16)


@pytest.mark.parametrize("m,n,k", [(128, 128, 64), (256, 384, 512), (64, 1000, 2048), (200, 136, 72),
                                   (1024, 2048, 512), (64, 4096, 25088)])
def test_gemm_nt_matches_fp32(cuda_device, m, n, k):
    torch.manual_seed(m + n + k)
    a = _bf16(torch.randn(m, k, device=cuda_device))
    b = _bf16(torch.randn(n, k, device=cuda_device) / (k ** 0.5))
    got = native.gemm_nt(a, b, out_dtype=torch.float32)
    ref = a.float() @ b.float().t()
    torch.cuda.synchronize()
    # fp32 output of bf16 operands against the fp32 product of the same operands: only the summation order
    # differs, so a mis-addressed k-slice (1/32 of K and up) cannot hide inside the tolerance
    assert torch.allclose(got, ref, rtol=1e-3, atol=1e-3), (got - ref).abs().max().item()


@pytest.mark.parametrize("act", [native.ACT_NONE, native.ACT_RELU, native.ACT_GELU])
@pytest.mark.parametrize("bias_dtype", [torch.float32, torch.bfloat16])
def test_gemm_epilogue(cuda_device, act, bias_dtype):
    torch.manual_seed(7)
    m, n, k = 192, 320, 256
    a = _bf16(torch.randn(m, k, device=cuda_device))
    b = _bf16(torch.randn(n, k, device=cuda_device) / 16)
    bias = torch.randn(n, device=cuda_device).to(bias_dtype)
    got = native.gemm_nt(a, b, bias, act).float()
    ref = a.float() @ b.float().t() + bias.float()
    if act == native.ACT_RELU:
        ref = torch.relu(ref)
    elif act == native.ACT_GELU:
        ref = torch.nn.functional.gelu(ref, approximate="tanh")
    assert torch.allclose(got, ref, rtol=3e-2, atol=3e-2), (got - ref).abs().max().item()


@pytest.fixture
def pair_gemm():
    """Every eligible GEMM on the 2-CTA (cta_group::2) kernel, whatever its size."""
    native.lib().tb200_gemm_set_pair_mode(1)
    yield
    native.lib().tb200_gemm_set_pair_mode(-1)


@pytest.mark.parametrize("m,n,k", [(256, 256, 64), (512, 512, 1024), (300, 520, 200), (128, 264, 72), (1000, 1000, 520),
                                   (256, 4096, 25088), (2048, 1536, 4096), (64, 256, 8192)])
def test_pair_gemm_nt_matches_fp32(cuda_device, pair_gemm, m, n, k):
    """256 x 256 pair tiles (UMMA M = 256 over two SMs): full tiles, M / N / K tails, one tile, split-K."""
    torch.manual_seed(m + n + k)
    a = _bf16(torch.randn(m, k, device=cuda_device))
    b = _bf16(torch.randn(n, k, device=cuda_device) / (k ** 0.5))
    ref = a.float() @ b.float().t()
    got = native.gemm_nt(a, b, out_dtype=torch.float32)
    assert torch.allclose(got, ref, rtol=1e-3, atol=1e-3), (got - ref).abs().max().item()
    got16 = native.gemm_nt(a, b)
    assert torch.allclose(got16.float(), ref, rtol=1e-2, atol=1e-2), (got16.float() - ref).abs().max().item()
    native.lib().tb200_gemm_set_pair_mode(0)
    one = native.gemm_nt(a, b)
    native.lib().tb200_gemm_set_pair_mode(1)
    if native.lib().tb200_gemm_plan_splits(m, n, k) == 1 and k <= 1024:
        # same k order inside a tile, no split-K on either: the two kernels agree to the bit
        assert torch.equal(one, got16) or (one.float() - got16.float()).abs().max().item() < 2e-2


@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (True, True), (True, False)])
@pytest.mark.parametrize("m,n,k", [(512, 768, 320), (264, 1000, 136)])
def test_pair_gemm_mn_major(cuda_device, pair_gemm, a_mn, b_mn, m, n, k):
    torch.manual_seed(5)
    a = _bf16(torch.randn(m, k, device=cuda_device))
    b = _bf16(torch.randn(n, k, device=cuda_device) / (k ** 0.5))
    got = native.gemm(a.t().contiguous() if a_mn else a, b.t().contiguous() if b_mn else b, a_mn_major=a_mn,
                      b_mn_major=b_mn, out_dtype=torch.float32)
    ref = a.float() @ b.float().t()
    assert torch.allclose(got, ref, rtol=1e-3, atol=1e-3), (got - ref).abs().max().item()


@pytest.mark.parametrize("act", [native.ACT_NONE, native.ACT_RELU, native.ACT_GELU])
@pytest.mark.parametrize("bias_dtype", [torch.float32, torch.bfloat16])
def test_pair_gemm_epilogue(cuda_device, pair_gemm, act, bias_dtype):
    torch.manual_seed(7)
    m, n, k = 448, 576, 256
    a = _bf16(torch.randn(m, k, device=cuda_device))
    b = _bf16(torch.randn(n, k, device=cuda_device) / 16)
    bias = torch.randn(n, device=cuda_device).to(bias_dtype)
    got = native.gemm_nt(a, b, bias, act).float()
    ref = a.float() @ b.float().t() + bias.float()
    if act == native.ACT_RELU:
        ref = torch.relu(ref)
    elif act == native.ACT_GELU:
        ref = torch.nn.functional.gelu(ref, approximate="tanh")
    assert torch.allclose(got, ref, rtol=3e-2, atol=3e-2), (got - ref).abs().max().item()


def test_pair_gemm_many_tiles_repeatable(cuda_device, pair_gemm):
    """More tiles than pairs (persistent loop, both accumulator stages, stage-ring wrap) - twice, same bits."""
    a = _bf16(torch.randn(4096, 768, device=cuda_device))
    b = _bf16(torch.randn(5120, 768, device=cuda_device) / 27)
    x = native.gemm_nt(a, b)
    y = native.gemm_nt(a, b)
    ref = a.float() @ b.float().t()
    assert torch.equal(x, y)
    assert torch.allclose(x.float(), ref, rtol=1e-2, atol=1e-2)


def test_gemm_repeatable(cuda_device):
    a = _bf16(torch.randn(512, 1024, device=cuda_device))
    b = _bf16(torch.randn(768, 1024, device=cuda_device))
    x = native.gemm_nt(a, b)
    y = native.gemm_nt(a, b)
    assert torch.equal(x, y)


def test_transpose(cuda_device):
    x = _bf16(torch.randn(333, 517, device=cuda_device))
    assert torch.equal(native.transpose(x), x.t().contiguous())


def test_u8_norm(cuda_device):
    x = torch.randint(0, 256, (3, 17, 19, 3), dtype=torch.uint8, device=cuda_device)
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    got = native.u8_to_bf16_norm(x, mean, std).float()
    ref = (x.float() / 255.0 - torch.tensor(mean, device=cuda_device)) / torch.tensor(std, device=cuda_device)
    assert torch.allclose(got, ref, rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_sumsq(cuda_device, dtype):
    x = torch.randn(1_000_003, device=cuda_device).to(dtype)
    got = native.sumsq(x).item()
    ref = x.double().pow(2).sum().item()
    assert abs(got - ref) / ref < 1e-4


@pytest.mark.parametrize("gdt", [torch.float32, torch.bfloat16])
def test_sgd_momentum(cuda_device, gdt):
    n = 8 * 4099
    torch.manual_seed(3)
    master = torch.randn(n, device=cuda_device)
    mom = torch.randn(n, device=cuda_device) * 0.1
    grad = torch.randn(n, device=cuda_device).to(gdt)
    lr, mu, wd, scale = 0.05, 0.9, 1e-3, 0.5
    hyper = torch.tensor([lr, mu, wd, scale], device=cuda_device)
    ref_g = grad.float() * scale + wd * master
    ref_v = mu * mom + ref_g
    ref_m = master - lr * ref_v
    out = torch.empty(n, dtype=torch.bfloat16, device=cuda_device)
    native.sgd_momentum(master, mom, grad, out, hyper)
    assert torch.allclose(mom, ref_v, rtol=1e-5, atol=1e-6)
    assert torch.allclose(master, ref_m, rtol=1e-5, atol=1e-6)
    assert torch.equal(out, ref_m.to(torch.bfloat16)) or torch.allclose(out.float(), ref_m, rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_ps_push_pull_emulated_world(cuda_device, world):
    """All W ranks live on this one GPU; flags are pre-posted from the host so that no kernel
    ever waits on another launch (B200 profiling recipe). Checks reduce + SGD + broadcast."""
    from tiresias_b200.parallel.symm import SymmetricWorld
    lib = native.lib()
    n = 8 * world * 515
    ranks = SymmetricWorld.emulated(world, cuda_device)
    for r in ranks:
        r.alloc("param", n * 2)
        r.alloc("grad", n * 2)
        r.alloc("flags", native.FLAG_WORDS * 4)
    torch.manual_seed(11)
    grads = [torch.randn(n, device=cuda_device).to(torch.bfloat16) for _ in range(world)]
    p0 = torch.randn(n, device=cuda_device)
    for r, g in zip(ranks, grads):
        r.buffers["grad"].view(torch.bfloat16).copy_(g)
        r.buffers["param"].view(torch.bfloat16).copy_(p0)
    lr, mu, wd = 0.1, 0.9, 1e-2
    hyper = torch.tensor([lr, mu, wd, 1.0 / world], device=cuda_device)
    epoch = torch.ones(1, dtype=torch.int32, device=cuda_device)
    status = torch.zeros(1, dtype=torch.int32, device=cuda_device)
    shard = n // world
    masters = [p0[r * shard:(r + 1) * shard].clone() for r in range(world)]
    moms = [torch.zeros(shard, device=cuda_device) for _ in range(world)]
    # stand in for the peers: everybody's gradients are final for epoch 1
    for r in ranks:
        r.post_flags("flags", native.flag_grad_ready(0, 0), world, 1)
    for r in ranks:
        w = r.native_world("grad", "param", "flags")
        counter = torch.zeros(1, dtype=torch.int32, device=cuda_device)
        native.check(lib.tb200_ps_push_pull(
            C.byref(w), 0, r.rank * shard, shard, native.BF16, masters[r.rank].data_ptr(),
            moms[r.rank].data_ptr(), hyper.data_ptr(), epoch.data_ptr(), counter.data_ptr(),
            status.data_ptr(), 1_000_000, 8, 1, native.stream_ptr()))
        torch.cuda.synchronize()
    assert status.item() == 0
    gsum = torch.zeros(n, device=cuda_device)
    for g in grads:                      # same rank order as the kernel
        gsum += g.float()
    ref_g = gsum / world + wd * p0
    ref_p = p0 - lr * ref_g              # momentum buffer starts at 0
    for r in ranks:
        got = r.buffers["param"].view(torch.bfloat16).float()
        assert torch.allclose(got, ref_p, rtol=1e-2, atol=1e-2)
        assert torch.equal(r.buffers["param"].view(torch.bfloat16), ranks[0].buffers["param"].view(torch.bfloat16))
    full_master = torch.cat(masters)
    assert torch.allclose(full_master, ref_p, rtol=1e-5, atol=1e-5)
    # PARAM_DONE flags were raised on every rank by every PS
    fl = ranks[0].buffers["flags"].view(torch.int32)
    assert all(int(fl[native.flag_param_done(0, s)]) == 1 for s in range(world))
    for r in ranks:
        r.close()


def test_push_timeout_reports_status(cuda_device):
    """A missing peer must turn into status != 0 (watchdog), not a hung GPU."""
    from tiresias_b200.parallel.symm import SymmetricWorld
    lib = native.lib()
    ranks = SymmetricWorld.emulated(2, cuda_device)
    n = 8 * 2 * 64
    for r in ranks:
        r.alloc("param", n * 2); r.alloc("grad", n * 2); r.alloc("flags", native.FLAG_WORDS * 4)
    hyper = torch.tensor([0.1, 0.9, 0.0, 0.5], device=cuda_device)
    epoch = torch.ones(1, dtype=torch.int32, device=cuda_device)
    status = torch.zeros(1, dtype=torch.int32, device=cuda_device)
    counter = torch.zeros(1, dtype=torch.int32, device=cuda_device)
    m = torch.zeros(n // 2, device=cuda_device); v = torch.zeros(n // 2, device=cuda_device)
    w = ranks[0].native_world("grad", "param", "flags")
    native.check(lib.tb200_ps_push_pull(C.byref(w), 0, 0, n // 2, native.BF16, m.data_ptr(), v.data_ptr(),
                                        hyper.data_ptr(), epoch.data_ptr(), counter.data_ptr(),
                                        status.data_ptr(), 2000, 4, 1, native.stream_ptr()))
    torch.cuda.synchronize()
    assert status.item() == 1
    for r in ranks:
        r.close()


@pytest.mark.parametrize("world", [1, 2, 4])
def test_direct_allreduce_emulated(cuda_device, world):
    from tiresias_b200.parallel.symm import SymmetricWorld
    lib = native.lib()
    n = 8 * world * 301
    ranks = SymmetricWorld.emulated(world, cuda_device)
    for r in ranks:
        r.alloc("buf", n * 4); r.alloc("flags", native.FLAG_WORDS * 4)
    torch.manual_seed(5)
    xs = [torch.randn(n, device=cuda_device) for _ in range(world)]
    for r, x in zip(ranks, xs):
        r.buffers["buf"].view(torch.float32).copy_(x)
    status = torch.zeros(1, dtype=torch.int32, device=cuda_device)
    slot = native.MAX_BUCKETS - 1
    for r in ranks:     # all contributions are in place, and stand in for the peers' DONE flags
        r.post_flags("flags", native.flag_grad_ready(slot, 0), world, 1)
        r.post_flags("flags", native.flag_param_done(slot, 0), world, 1)
    for r in ranks:
        w = r.native_world(None, None, "flags")
        ptrs = (C.c_void_p * native.MAX_WORLD)(*[q.buffers["buf"].ptrs[q.rank] for q in ranks])
        seq = torch.zeros(2, dtype=torch.int32, device=cuda_device)
        native.check(lib.tb200_allreduce(C.byref(w), ptrs, n, native.F32, 1.0 / world, 1, seq.data_ptr(),
                                         status.data_ptr(), 1_000_000, 8, native.stream_ptr()))
        torch.cuda.synchronize()
    assert status.item() == 0
    ref = sum(xs) / world
    # rank r reduced slice r and stored it everywhere
    for r in ranks:
        got = r.buffers["buf"].view(torch.float32)
        assert torch.allclose(got, ref, rtol=1e-5, atol=1e-5)
    for r in ranks:
        r.close()


def test_u8_norm_pad8(cuda_device):
    x = torch.randint(0, 256, (3, 9, 7, 3), dtype=torch.uint8, device=cuda_device)      # 189 pixels: tail of 1
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    got = native.u8_to_bf16_norm_pad8(x, mean, std)
    assert got.shape == (3, 9, 7, 8) and float(got[..., 3:].abs().sum()) == 0.0
    want = (x.float() / 255.0 - torch.tensor(mean, device=cuda_device)) / torch.tensor(std, device=cuda_device)
    assert torch.allclose(got[..., :3].float(), want, rtol=1e-2, atol=2e-2)


def test_pack_bf16_gathers_many_tensors(cuda_device):
    import ctypes as C
    torch.manual_seed(3)
    sizes = [5, 8, 64, 1000, 8192, 8193, 70000] + [17] * 80            # > TB200_PACK_MAX items, odd sizes
    srcs = [torch.randn(n, device=cuda_device).to(torch.bfloat16) for n in sizes]
    offs, total = [], 0
    for n in sizes:
        total = (total + 63) // 64 * 64
        offs.append(total)
        total += n
    flat = torch.zeros(total + 64, dtype=torch.bfloat16, device=cuda_device)
    items = (native.PackItem * len(sizes))(*[native.PackItem(s.data_ptr(), flat.data_ptr() + 2 * o, n)
                                             for s, o, n in zip(srcs, offs, sizes)])
    native.check(native.lib().tb200_pack_bf16(items, len(sizes), native.stream_ptr()))
    torch.cuda.synchronize()
    for s, o, n in zip(srcs, offs, sizes):
        assert torch.equal(flat[o:o + n], s)
    assert float(flat[offs[0] + sizes[0]:offs[1]].abs().sum()) == 0.0             # gaps untouched
    # an unaligned destination takes the scalar path
    dst = torch.zeros(100, dtype=torch.bfloat16, device=cuda_device)
    one = (native.PackItem * 1)(native.PackItem(srcs[3].data_ptr(), dst.data_ptr() + 2, 50))
    native.check(native.lib().tb200_pack_bf16(one, 1, native.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.equal(dst[1:51], srcs[3][:50]) and float(dst[0]) == 0.0


@pytest.mark.parametrize("m,n,k", [(128, 128, 64), (256, 384, 512), (200, 136, 72), (1024, 2048, 512), (8192, 512, 2048),
                                   (64, 1000, 2048)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (True, True), (True, False)])
def test_gemm_operand_layouts_match_fp32(cuda_device, m, n, k, a_mn, b_mn):
    if (a_mn and m % 8) or (b_mn and n % 8):
        pytest.skip("MN-major operands need a contiguous dimension that is a multiple of 8")
    g = torch.Generator(device=cuda_device).manual_seed(m + n + k)
    a = torch.randn(m, k, device=cuda_device, generator=g).to(torch.bfloat16)
    b = (torch.randn(n, k, device=cuda_device, generator=g) / k ** 0.5).to(torch.bfloat16)
    a_arg = a.t().contiguous() if a_mn else a
    b_arg = b.t().contiguous() if b_mn else b
    got = native.gemm(a_arg, b_arg, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.float32)
    ref = a.float() @ b.float().t()
    assert torch.allclose(got, ref, rtol=1e-3, atol=1e-3), (got - ref).abs().max().item()


# Single-GPU check of the pull GEMM (K2: B tiles TMA-loaded from the shard owners' memory) with emulated owners;
# the multi-rank trainers (tests/multi_worker.py, 2/4/8 GPUs) cover the same kernel over real peers.
@pytest.mark.parametrize("world", [2, 4])
def test_pull_gemm_emulated_owners(cuda_device, world):
    from tiresias_b200.parallel.symm import SymmetricWorld
    lib = native.lib()
    m, n, k = 96, 128 * world, 256                        # every owner holds 128 rows of B
    ranks = SymmetricWorld.emulated(world, cuda_device)
    for r in ranks:
        r.alloc("param", n * k * 2)
        r.alloc("grad", 64)
        r.alloc("flags", native.FLAG_WORDS * 4)
    torch.manual_seed(3)
    a = torch.randn(m, k, device=cuda_device).to(torch.bfloat16)
    b = (torch.randn(n, k, device=cuda_device) / k ** 0.5).to(torch.bfloat16)
    rows = n // world
    for r in ranks:                                        # owner r has only ITS shard up to date
        buf = r.buffers["param"].view(torch.bfloat16).view(n, k)
        buf.zero_()
        buf[r.rank * rows:(r.rank + 1) * rows].copy_(b[r.rank * rows:(r.rank + 1) * rows])
    epoch = torch.full((1,), 2, dtype=torch.int32, device=cuda_device)      # shards were written in step 1
    status = torch.zeros(1, dtype=torch.int32, device=cuda_device)
    me = ranks[0]
    me.post_flags("flags", native.flag_param_done(0, 0), world, 1)
    w = me.native_world("grad", "param", "flags")
    y = torch.empty(m, n, dtype=torch.bfloat16, device=cuda_device)
    sync = torch.zeros(native.PULL_SYNC_WORDS, dtype=torch.int32, device=cuda_device)
    for _ in range(2):                                      # twice: the arrival counters reset themselves
        native.check(lib.tb200_gemm_bf16_nt_pull(C.byref(w), 0, 0, rows, a.data_ptr(), y.data_ptr(), None, 0, m, n, k,
                                                 native.BF16, 0, epoch.data_ptr(), status.data_ptr(), 1_000_000,
                                                 sync.data_ptr(), native.stream_ptr()))
    torch.cuda.synchronize()
    assert status.item() == 0
    ref = a.float() @ b.float().t()
    assert torch.allclose(y.float(), ref, rtol=2e-2, atol=2e-2)
    # the copy engine left the full matrix in rank 0's own buffer
    assert torch.equal(me.buffers["param"].view(torch.bfloat16).view(n, k), b)
    for r in ranks:
        r.close()
