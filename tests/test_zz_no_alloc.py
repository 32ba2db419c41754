"""The allocation-free GEMM entry point, driven through the C ABI exactly as a C host would.

Jobs that share a GPU with other jobs (an allocation synchronises the whole device and can fail
on a fully subscribed GPU) and C hosts without an allocator call ``tb200_gemm_bf16_nt_noalloc``
instead of ``tb200_gemm_bf16_nt``. Its split-K partial sums live in a scratch array built into the
library, so this file checks that (1) it returns what the allocating entry point returns, on two
back-to-back calls that reuse the shared scratch with different shapes, (2) the device's free
memory does not move across a call, and (3) a problem whose partial sums do not fit the built-in
scratch is refused with a message instead of being launched.

Collected last on purpose (``test_zz_``): a run that stops at the first failure has then already
exercised everything else. It loads the library itself, so it also passes alone in a fresh process.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 16 << 20      # TB200_GEMM_NOALLOC_SCRATCH_BYTES in include/tiresias_b200.h
BF16, F32 = 1, 0


@pytest.fixture(scope="module")
def cabi(cuda_device):
    """The shared library opened with nothing but ctypes and the declarations of the header."""
    from tiresias_b200.ops.native import LIB_PATH
    lib = C.CDLL(str(LIB_PATH))
    sig = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
           C.c_int, C.c_int, C.c_void_p]
    for name in ("tb200_gemm_bf16_nt", "tb200_gemm_bf16_nt_noalloc"):
        fn = getattr(lib, name)
        fn.argtypes, fn.restype = sig, C.c_int
    lib.tb200_last_error.restype = C.c_char_p
    lib.tb200_gemm_plan_splits.argtypes, lib.tb200_gemm_plan_splits.restype = [C.c_int] * 3, C.c_int
    return lib


def _problem(device, m, n, k, seed):
    g = torch.Generator(device=device).manual_seed(seed)
    a = torch.randn(m, k, device=device, generator=g).to(torch.bfloat16)
    b = (torch.randn(n, k, device=device, generator=g) / k ** 0.5).to(torch.bfloat16)
    bias = torch.randn(n, device=device, generator=g)
    return a, b, bias


def _call(lib, fn_name, a, b, bias, out):
    m, k = a.shape
    n = b.shape[0]
    stream = torch.cuda.current_stream().cuda_stream
    rc = getattr(lib, fn_name)(a.data_ptr(), b.data_ptr(), out.data_ptr(), bias.data_ptr(), F32,
                               m, n, k, F32, 1, stream)
    assert rc == 0, lib.tb200_last_error().decode()


def test_noalloc_matches_allocating_entry_point_twice_in_a_row(cabi, cuda_device):
    # first a split-K problem whose partial sums take several MiB of the scratch, then a smaller split-K
    # problem that finds the first call's partial sums still lying in the shared array
    shapes = [(64, 4096, 25088), (64, 1000, 2048)]
    # (64 x 4096 x 25088 plans 9 splits = 9 MiB of fp32 partials on a 148-SM B200)
    problems = [_problem(cuda_device, *s, seed=i) for i, s in enumerate(shapes)]
    wants = []
    for a, b, bias in problems:
        want = torch.empty(a.shape[0], b.shape[0], device=cuda_device)
        _call(cabi, "tb200_gemm_bf16_nt", a, b, bias, want)
        wants.append(want)
    outs = [torch.full_like(w, float("nan")) for w in wants]
    torch.cuda.synchronize()

    # set-up call (resolves the scratch address on this device), then measure across both calls
    _call(cabi, "tb200_gemm_bf16_nt_noalloc", *problems[1], outs[1])
    torch.cuda.synchronize()
    outs[1].fill_(float("nan"))
    free_before, _ = torch.cuda.mem_get_info()
    for (a, b, bias), out in zip(problems, outs):        # two calls in a row, no sync in between
        _call(cabi, "tb200_gemm_bf16_nt_noalloc", a, b, bias, out)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after >= free_before, "the allocation-free entry point took device memory"

    for (a, b, bias), got, want in zip(problems, outs, wants):
        ref = torch.relu(a.float() @ b.float().t() + bias)
        # same plan, same kernels, same summation order: the two entry points agree to the bit; against an
        # fp32 product of the same bf16 operands only the accumulation order differs
        assert torch.equal(got, want), (got - want).abs().max().item()
        assert torch.allclose(got, ref, rtol=2e-3, atol=2e-3), (got - ref).abs().max().item()


def test_noalloc_without_split_uses_no_scratch(cabi, cuda_device):
    a, b, bias = _problem(cuda_device, 2048, 4096, 512, seed=5)     # 16 x 16 tiles: no split-K
    want = torch.empty(2048, 4096, device=cuda_device)
    got = torch.empty_like(want)
    _call(cabi, "tb200_gemm_bf16_nt", a, b, bias, want)
    _call(cabi, "tb200_gemm_bf16_nt_noalloc", a, b, bias, got)
    torch.cuda.synchronize()
    assert cabi.tb200_gemm_plan_splits(2048, 4096, 512) == 1
    assert torch.equal(got, want)


def test_noalloc_rejects_problems_beyond_the_scratch(cabi, cuda_device):
    m, n, k = 1024, 2048, 8192            # 64 tiles of 128 x 256 on 148 SMs -> split along K
    splits = cabi.tb200_gemm_plan_splits(m, n, k)          # 16 on a 148-SM B200 = 128 MiB of partials
    a, b, bias = _problem(cuda_device, m, n, k, seed=9)
    out = torch.zeros(m, n, device=cuda_device)
    stream = torch.cuda.current_stream().cuda_stream
    rc = cabi.tb200_gemm_bf16_nt_noalloc(a.data_ptr(), b.data_ptr(), out.data_ptr(), bias.data_ptr(),
                                         F32, m, n, k, F32, 0, stream)
    torch.cuda.synchronize()
    assert (rc == 0) == (splits * m * n * 4 <= SCRATCH_BYTES), (rc, splits)
    if rc == 0:
        pytest.skip("this GPU's SM count plans the problem without exceeding the scratch")
    msg = cabi.tb200_last_error().decode()
    assert "scratch" in msg and str(SCRATCH_BYTES) in msg, msg
    assert float(out.abs().sum()) == 0.0, "a refused call must not launch anything"


def test_noalloc_on_a_second_device_of_the_same_process(cabi, cuda_device):
    """Kernel attributes and the scratch address are per-device state: a host that drives two GPUs from
    one process must get the same answer on the second GPU (skipped on a one-GPU box)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    outs = []
    for dev in (0, 1):
        with torch.cuda.device(dev):
            d = torch.device("cuda", dev)
            a, b, bias = _problem(d, 64, 1000, 2048, seed=3)
            got, want = torch.empty(64, 1000, device=d), torch.empty(64, 1000, device=d)
            _call(cabi, "tb200_gemm_bf16_nt_noalloc", a, b, bias, got)
            _call(cabi, "tb200_gemm_bf16_nt", a, b, bias, want)
            torch.cuda.synchronize(d)
            assert torch.equal(got, want)
            outs.append(got.cpu())
    assert torch.equal(outs[0], outs[1])
