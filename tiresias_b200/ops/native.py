"""ctypes bindings for ``libtiresias_b200.so`` (C ABI in ``include/tiresias_b200.h``).

The library is built in-tree by :mod:`tiresias_b200.build`. On a machine with a GPU a missing
or unloadable library is an error (no silent PyTorch fallback); on the CPU-only build box the
module imports fine and :func:`lib` raises only when a kernel is actually requested.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional, Sequence

import torch

MAX_WORLD = 8
MAX_BUCKETS = 256
MAX_RING_CTAS = 64
IPC_HANDLE_BYTES = 64
FLAG_WORDS = 2 * MAX_BUCKETS * MAX_WORLD + 2 * MAX_RING_CTAS
PULL_SYNC_WORDS = 4096        # TB200_PULL_SYNC_WORDS
BN_SCRATCH_FLOATS = 1024 + 2 * 64 * 1024 + 4 * 4096 + 2 * 1024 * 1024   # TB200_BN_SCRATCH_FLOATS

F32, BF16 = 0, 1
ACT_NONE, ACT_RELU, ACT_GELU = 0, 1, 2

LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libtiresias_b200.so"


def flag_grad_ready(bucket: int, src: int) -> int:
    return bucket * MAX_WORLD + src


def flag_param_done(bucket: int, src: int) -> int:
    return MAX_BUCKETS * MAX_WORLD + bucket * MAX_WORLD + src


class World(C.Structure):
    """Mirror of ``tb200_world_t``."""

    _fields_ = [
        ("world", C.c_int),
        ("rank", C.c_int),
        ("grad", C.c_void_p * MAX_WORLD),
        ("param", C.c_void_p * MAX_WORLD),
        ("flags", C.c_void_p * MAX_WORLD),
        ("grad_mc", C.c_void_p),
        ("param_mc", C.c_void_p),
    ]


class PackItem(C.Structure):
    """Mirror of ``tb200_pack_item_t``."""

    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("n", C.c_ulonglong)]


class SgdHyper(C.Structure):
    _fields_ = [("lr", C.c_float), ("momentum", C.c_float), ("weight_decay", C.c_float),
                ("grad_scale", C.c_float)]


_lib: Optional[C.CDLL] = None


class NativeError(RuntimeError):
    pass


def _declare(lib: C.CDLL) -> None:
    vp, sz, u32p, ip, st = C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p
    lib.tb200_version.restype = C.c_int
    lib.tb200_last_error.restype = C.c_char_p
    lib.tb200_launch_count.restype = C.c_ulonglong
    lib.tb200_launch_count_reset.restype = None
    lib.tb200_symm_alloc.argtypes = [sz, C.POINTER(vp)]
    lib.tb200_symm_free.argtypes = [vp]
    lib.tb200_ipc_export.argtypes = [vp, C.c_char_p]
    lib.tb200_ipc_open.argtypes = [C.c_char_p, C.POINTER(vp)]
    lib.tb200_ipc_close.argtypes = [vp]
    lib.tb200_enable_peer_access.argtypes = [C.c_int, C.c_int]
    lib.tb200_flags_fill.argtypes = [u32p, sz, sz, C.c_uint32, st]
    lib.tb200_sgd_momentum.argtypes = [vp, vp, vp, C.c_int, vp, C.c_int, sz, vp, st]
    lib.tb200_ps_push_pull.argtypes = [C.POINTER(World), C.c_int, sz, sz, C.c_int, vp, vp, vp, u32p,
                                       u32p, ip, C.c_ulonglong, C.c_int, C.c_int, st]
    lib.tb200_step_end.argtypes = [C.POINTER(World), C.c_int, u32p, ip, C.c_ulonglong, st]
    lib.tb200_allreduce.argtypes = [C.POINTER(World), C.POINTER(vp), sz, C.c_int, C.c_float, C.c_int,
                                    u32p, ip, C.c_ulonglong, C.c_int, st]
    lib.tb200_gemm_bf16_nt.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_int, st]
    lib.tb200_gemm_bf16.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_int, C.c_int, st]
    lib.tb200_gemm_bf16.restype = C.c_int
    ci = C.c_int
    lib.tb200_conv2d_fprop.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, ci, ci, st]
    lib.tb200_conv2d_fprop.restype = ci
    lib.tb200_conv2d_dgrad.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, ci, ci, st]
    lib.tb200_conv2d_dgrad.restype = ci
    lib.tb200_conv2d_wgrad.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, ci, ci, ci, st]
    lib.tb200_conv2d_wgrad.restype = ci
    ull, ullp, vpp, ip_ = C.c_ulonglong, C.POINTER(C.c_ulonglong), C.POINTER(vp), C.POINTER(C.c_int)
    lib.tb200_vmm_round.argtypes = [ci, sz, C.POINTER(sz)]
    lib.tb200_vmm_alloc.argtypes = [sz, ullp, vpp, ip_]
    lib.tb200_vmm_import.argtypes = [ci, sz, ullp, vpp]
    lib.tb200_vmm_release.argtypes = [ull, vp, sz]
    lib.tb200_mc_create.argtypes = [ci, sz, ullp, ip_]
    lib.tb200_mc_import.argtypes = [ci, ullp]
    lib.tb200_mc_add_device.argtypes = [ull]
    lib.tb200_mc_bind.argtypes = [ull, ull, sz, vpp]
    lib.tb200_mc_release.argtypes = [ull, vp, sz]
    for fn in ("tb200_vmm_round", "tb200_vmm_alloc", "tb200_vmm_import", "tb200_vmm_release", "tb200_mc_create",
               "tb200_mc_import", "tb200_mc_add_device", "tb200_mc_bind", "tb200_mc_release"):
        getattr(lib, fn).restype = ci
    lib.tb200_conv2d_fprop_halo.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, st]
    lib.tb200_conv2d_fprop_halo.restype = ci
    lib.tb200_stem7x7_pack.argtypes = [vp, vp, ci, ci, ci, C.POINTER(C.c_float), C.POINTER(C.c_float), st]
    lib.tb200_stem7x7_pack.restype = ci
    lib.tb200_stem7x7_weight_pack.argtypes = [vp, vp, ci, st]
    lib.tb200_stem7x7_weight_pack.restype = ci
    lib.tb200_stem7x7_wgrad_unpack.argtypes = [vp, vp, ci, ci, st]
    lib.tb200_stem7x7_wgrad_unpack.restype = ci
    lib.tb200_gemm_bf16_nt_stats.argtypes = [vp, vp, vp, ci, ci, ci, vp, sz, ip_, st]
    lib.tb200_gemm_bf16_nt_stats.restype = ci
    lib.tb200_conv2d_fprop_stats.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, ci, ci, vp, sz, ip_, st]
    lib.tb200_conv2d_fprop_stats.restype = ci
    lib.tb200_conv_stats_floats.argtypes = [ci]
    lib.tb200_conv_stats_floats.restype = sz
    lib.tb200_conv3x3_fprop.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, vp, sz, ip_, st]
    lib.tb200_conv3x3_fprop.restype = ci
    lib.tb200_conv3x3_dgrad.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, st]
    lib.tb200_conv3x3_dgrad.restype = ci
    lib.tb200_conv3x3_fprop_bias.argtypes = [vp, vp, vp, ci, ci, vp, ci, ci, ci, ci, ci, st]
    lib.tb200_conv3x3_fprop_bias.restype = ci
    lib.tb200_conv3x3_supported.argtypes = [ci, ci, ci, ci, ci, ci]
    lib.tb200_conv3x3_supported.restype = ci
    lib.tb200_conv3x3_wgrad_workspace.argtypes = [ci, ci, ci, ci, ci]
    lib.tb200_conv3x3_wgrad_workspace.restype = sz
    lib.tb200_conv3x3_wgrad.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, vp, sz, st]
    lib.tb200_conv3x3_wgrad.restype = ci
    lib.tb200_bn_fwd_train_stats.argtypes = [vp, vp, vp, vp, vp, ci, vp, vp, C.c_float, C.c_float, vp, vp, vp, vp, ci,
                                             sz, ci, ci, st]
    lib.tb200_bn_fwd_train_stats.restype = ci
    lib.tb200_bn_mask_supported.argtypes = [sz, ci]
    lib.tb200_bn_mask_supported.restype = ci
    lib.tb200_bn_fwd_train_ex.argtypes = [vp, vp, vp, vp, vp, ci, vp, vp, C.c_float, C.c_float, vp, vp, vp, vp, ci,
                                          vp, sz, ci, ci, st]
    lib.tb200_bn_fwd_train_ex.restype = ci
    lib.tb200_bn_bwd_ex.argtypes = [vp, vp, vp, vp, vp, vp, ci, vp, vp, vp, vp, vp, vp, vp, sz, ci, ci, st]
    lib.tb200_bn_bwd_ex.restype = ci
    lib.tb200_bn_finalize.argtypes = [vp, ci, C.c_double, ci, vp, vp, ci, vp, vp, C.c_float, C.c_float, vp, vp, vp, st]
    lib.tb200_bn_finalize.restype = ci
    lib.tb200_nvls_supported.argtypes = [ci]
    lib.tb200_nvls_supported.restype = ci
    lib.tb200_nvls_selftest.argtypes = [ci, sz, ci, C.c_char_p, sz]
    lib.tb200_nvls_selftest.restype = ci
    lib.tb200_maxpool_fwd.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, st]
    lib.tb200_maxpool_fwd.restype = ci
    lib.tb200_maxpool_bwd.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, st]
    lib.tb200_maxpool_bwd.restype = ci
    lib.tb200_attention_fwd.argtypes = [vp, vp, vp, ci, ci, ci, ci, st]
    lib.tb200_attention_fwd.restype = ci
    lib.tb200_attention_bwd.argtypes = [vp, vp, vp, vp, vp, vp, ci, ci, ci, ci, st]
    lib.tb200_attention_bwd.restype = ci
    lib.tb200_layernorm_scratch_floats.argtypes = [ci]
    lib.tb200_layernorm_scratch_floats.restype = sz
    lib.tb200_layernorm_fwd.argtypes = [vp, vp, vp, vp, ci, vp, vp, sz, ci, C.c_float, st]
    lib.tb200_layernorm_fwd.restype = ci
    lib.tb200_layernorm_bwd.argtypes = [vp, vp, vp, ci, vp, vp, vp, vp, vp, vp, sz, sz, ci, st]
    lib.tb200_layernorm_bwd.restype = ci
    lib.tb200_gemm_bf16_add.argtypes = [vp, vp, vp, vp, ci, ci, ci, ci, ci, st]
    lib.tb200_gemm_bf16_add.restype = ci
    lib.tb200_gemm_set_pair_mode.argtypes = [ci]
    lib.tb200_gemm_set_pair_mode.restype = None
    lib.tb200_gemm_release_stream.argtypes = [st]
    lib.tb200_gemm_release_stream.restype = ci
    lib.tb200_gemm_reserve.argtypes = [sz]
    lib.tb200_gemm_bf16_nt_pull.argtypes = [C.POINTER(World), C.c_int, sz, C.c_int, vp, vp, vp,
                                            C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                            u32p, ip,
                                            C.c_ulonglong, vp, st]
    lib.tb200_u8_to_bf16_norm.argtypes = [vp, vp, sz, C.POINTER(C.c_float), C.POINTER(C.c_float), st]
    lib.tb200_u8_to_bf16_norm_pad8.argtypes = [vp, vp, sz, C.POINTER(C.c_float), C.POINTER(C.c_float), st]
    lib.tb200_u8_to_bf16_norm_pad8.restype = C.c_int
    lib.tb200_transpose_bf16.argtypes = [vp, vp, C.c_int, C.c_int, st]
    lib.tb200_maxpool3x3s2_fwd.argtypes = [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, st]
    lib.tb200_maxpool3x3s2_fwd.restype = C.c_int
    lib.tb200_maxpool3x3s2_bwd.argtypes = [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, st]
    lib.tb200_maxpool3x3s2_bwd.restype = C.c_int
    lib.tb200_pack_bf16.argtypes = [C.POINTER(PackItem), C.c_int, st]
    lib.tb200_pack_bf16.restype = C.c_int
    lib.tb200_sumsq_scratch_floats.restype = C.c_size_t
    lib.tb200_sumsq.argtypes = [vp, C.c_int, sz, vp, vp, st]
    lib.tb200_bn_scratch_floats.restype = C.c_size_t
    lib.tb200_bn_fwd_train.argtypes = [vp, vp, vp, vp, vp, C.c_int, vp, vp, C.c_float, C.c_float, vp, vp, vp,
                                       sz, C.c_int, C.c_int, st]
    lib.tb200_bn_apply.argtypes = [vp, vp, vp, vp, sz, C.c_int, C.c_int, st]
    lib.tb200_bn_bwd.argtypes = [vp, vp, vp, vp, vp, C.c_int, vp, vp, vp, vp, vp, vp, vp, sz, C.c_int, C.c_int, st]
    for name in ("tb200_symm_alloc", "tb200_symm_free", "tb200_ipc_export", "tb200_ipc_open",
                 "tb200_ipc_close", "tb200_enable_peer_access", "tb200_flags_fill",
                 "tb200_sgd_momentum", "tb200_ps_push_pull", "tb200_step_end", "tb200_allreduce",
                 "tb200_gemm_bf16_nt", "tb200_gemm_reserve", "tb200_gemm_bf16_nt_pull",
                 "tb200_u8_to_bf16_norm", "tb200_transpose_bf16", "tb200_sumsq", "tb200_bn_fwd_train",
                 "tb200_bn_apply", "tb200_bn_bwd"):
        getattr(lib, name).restype = C.c_int


def available() -> bool:
    return LIB_PATH.exists()


def lib() -> C.CDLL:
    """Load (once) and return the native library; raises :class:`NativeError` if absent."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise NativeError(
                f"{LIB_PATH} is missing: run `python -m tiresias_b200.build` (nvcc, sm_100a). "
                "There is no PyTorch fallback for the native kernels.")
        # torch has already loaded libcudart; make sure ours resolves to the same runtime.
        _lib = C.CDLL(str(LIB_PATH), mode=os.RTLD_GLOBAL if hasattr(os, "RTLD_GLOBAL") else 0)
        _declare(_lib)
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise NativeError(lib().tb200_last_error().decode(errors="replace"))


def stream_ptr(stream: Optional[torch.cuda.Stream] = None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float32:
        return F32
    if dt == torch.bfloat16:
        return BF16
    raise TypeError(f"unsupported dtype {dt}")


def launch_count() -> int:
    return int(lib().tb200_launch_count())


def launch_count_reset() -> None:
    lib().tb200_launch_count_reset()


# ---- thin tensor-level wrappers ----------------------------------------------------------------

def sgd_momentum(master: torch.Tensor, momentum: torch.Tensor, grad: torch.Tensor,
                 param_out: torch.Tensor, hyper_dev: torch.Tensor) -> None:
    """K4: fused SGD+momentum+weight-decay step. ``hyper_dev``: 4 fp32 on device
    (lr, momentum, weight_decay, grad_scale)."""
    n = master.numel()
    assert momentum.numel() == n and grad.numel() == n and param_out.numel() == n
    check(lib().tb200_sgd_momentum(master.data_ptr(), momentum.data_ptr(), grad.data_ptr(),
                                   dtype_code(grad.dtype), param_out.data_ptr(),
                                   dtype_code(param_out.dtype), n, hyper_dev.data_ptr(), stream_ptr()))


def gemm_nt(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None,
            act: int = ACT_NONE, out_dtype: torch.dtype = torch.bfloat16,
            out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """K5: ``act(a @ b.T + bias)`` on the tcgen05 tensor cores. a [M,K], b [N,K], bf16."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert a.is_contiguous() and b.is_contiguous() and a.shape[1] == b.shape[1]
    m, k = a.shape
    n = b.shape[0]
    if out is None:
        out = torch.empty((m, n), dtype=out_dtype, device=a.device)
    bias_ptr, bias_dt = 0, F32
    if bias is not None:
        assert bias.numel() == n and bias.is_contiguous()
        bias_ptr, bias_dt = bias.data_ptr(), dtype_code(bias.dtype)
    check(lib().tb200_gemm_bf16_nt(a.data_ptr(), b.data_ptr(), out.data_ptr(), bias_ptr, bias_dt,
                                   m, n, k, dtype_code(out.dtype), act, stream_ptr()))
    return out


def gemm_nt_stats(a: torch.Tensor, b: torch.Tensor):
    """``a @ b.T`` (bf16 out) plus the per-channel (sum, sum of squares) partials of the output from the
    epilogue (tb200_gemm_bf16_nt_stats). Returns ``(out, ConvStats or None)`` — None when the problem had to be
    split along K (tiny batch) and carries no statistics."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert a.is_contiguous() and b.is_contiguous() and a.shape[1] == b.shape[1]
    m, k = a.shape
    n = b.shape[0]
    out = torch.empty((m, n), dtype=torch.bfloat16, device=a.device)
    buf = conv_stats_buffer(n, a.device)
    nparts = C.c_int(0)
    check(lib().tb200_gemm_bf16_nt_stats(a.data_ptr(), b.data_ptr(), out.data_ptr(), m, n, k, buf.data_ptr(), buf.numel(),
                                         C.byref(nparts), stream_ptr()))
    return out, (ConvStats(buf[: nparts.value], nparts.value, m) if nparts.value else None)


def gemm(a: torch.Tensor, b: torch.Tensor, a_mn_major: bool = False, b_mn_major: bool = False,
         bias: Optional[torch.Tensor] = None, act: int = ACT_NONE, out_dtype: torch.dtype = torch.bfloat16,
         out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """K5 with selectable operand layouts. ``a`` is [M,K] (or [K,M] when ``a_mn_major``), ``b`` is [N,K]
    (or [K,N] when ``b_mn_major``); returns ``act(A @ B^T + bias)`` as [M,N]."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_contiguous() and b.is_contiguous()
    k, m = (a.shape if a_mn_major else a.shape[::-1])
    kb, n = (b.shape if b_mn_major else b.shape[::-1])
    assert k == kb, (a.shape, b.shape)
    if out is None:
        out = torch.empty((m, n), dtype=out_dtype, device=a.device)
    bias_ptr, bias_dt = 0, F32
    if bias is not None:
        bias_ptr, bias_dt = bias.data_ptr(), dtype_code(bias.dtype)
    check(lib().tb200_gemm_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), bias_ptr, bias_dt, m, n, k,
                                int(a_mn_major), int(b_mn_major), dtype_code(out.dtype), act, stream_ptr()))
    return out


def gemm_add(a: torch.Tensor, b: torch.Tensor, addend: torch.Tensor, a_mn_major: bool = False,
             b_mn_major: bool = False, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``A @ B^T + addend`` (bf16 [M, N]) with the addend folded into the GEMM's epilogue; ``out`` may be the
    addend itself."""
    assert a.dtype == b.dtype == addend.dtype == torch.bfloat16
    assert a.is_contiguous() and b.is_contiguous() and addend.is_contiguous()
    k, m = (a.shape if a_mn_major else a.shape[::-1])
    kb, n = (b.shape if b_mn_major else b.shape[::-1])
    assert k == kb and tuple(addend.shape) == (m, n), (a.shape, b.shape, addend.shape)
    if out is None:
        out = torch.empty((m, n), dtype=torch.bfloat16, device=a.device)
    check(lib().tb200_gemm_bf16_add(a.data_ptr(), b.data_ptr(), out.data_ptr(), addend.data_ptr(), m, n, k,
                                    int(a_mn_major), int(b_mn_major), stream_ptr()))
    return out


def nvls_supported(device: int = 0) -> bool:
    """Can this GPU join an NVLS multicast object (NVSwitch fabric)?"""
    return lib().tb200_nvls_supported(device) == 1


def nvls_selftest(ndev: int, nbytes: int = 64 << 20, ctas: int = 148) -> Optional[dict]:
    """Checks and times multimem.ld_reduce / multimem.st / the one-shot NVLS all-reduce over GPUs
    0..ndev-1 of this process (csrc/nvls.cu). Returns the report, or None when the box has no NVLS."""
    import json
    buf = C.create_string_buffer(2048)
    rc = lib().tb200_nvls_selftest(ndev, nbytes, ctas, buf, len(buf))
    if rc == 2:
        return None
    if rc != 0 and not buf.value:
        check(rc)
    rep = json.loads(buf.value.decode())
    if rc != 0:
        rep["error"] = lib().tb200_last_error().decode(errors="replace")
    return rep


def _nhwc(t: torch.Tensor) -> torch.Tensor:
    """Logical NCHW tensor stored channels_last -> its contiguous [N,H,W,C] view."""
    v = t.permute(0, 2, 3, 1)
    assert v.is_contiguous(), "expected a channels_last tensor"
    return v


def conv2d_fprop_stats(x: torch.Tensor, weight: torch.Tensor, stride: int, pad: int):
    """:func:`conv2d_fprop` plus the epilogue's BatchNorm statistics: ``(y, ConvStats or None)``."""
    assert x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16
    xv, wv = _nhwc(x), _nhwc(weight)
    n, h, w, cin = xv.shape
    cout, r, s, _ = wv.shape
    oh, ow = (h + 2 * pad - r) // stride + 1, (w + 2 * pad - s) // stride + 1
    y = torch.empty((n, oh, ow, cout), dtype=torch.bfloat16, device=x.device)
    buf = conv_stats_buffer(cout, x.device)
    nparts = C.c_int(0)
    check(lib().tb200_conv2d_fprop_stats(xv.data_ptr(), wv.data_ptr(), y.data_ptr(), n, h, w, cin, cout, r, s, stride, pad,
                                         buf.data_ptr(), buf.numel(), C.byref(nparts), stream_ptr()))
    st = ConvStats(buf[: nparts.value], nparts.value, n * oh * ow) if nparts.value else None
    return y.permute(0, 3, 1, 2), st


def conv2d_fprop(x: torch.Tensor, weight: torch.Tensor, stride: int, pad: int) -> torch.Tensor:
    """Implicit-GEMM convolution forward (tb200_conv2d_fprop). ``x`` [N,Cin,H,W] and ``weight``
    [Cout,Cin,R,S] are bf16 channels_last; returns y [N,Cout,OH,OW] channels_last."""
    assert x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16
    xv, wv = _nhwc(x), _nhwc(weight)
    n, h, w, cin = xv.shape
    cout, r, s, _ = wv.shape
    oh, ow = (h + 2 * pad - r) // stride + 1, (w + 2 * pad - s) // stride + 1
    y = torch.empty((n, oh, ow, cout), dtype=torch.bfloat16, device=x.device)
    check(lib().tb200_conv2d_fprop(xv.data_ptr(), wv.data_ptr(), y.data_ptr(), n, h, w, cin, cout, r, s,
                                   stride, pad, stream_ptr()))
    return y.permute(0, 3, 1, 2)


def conv2d_fprop_halo(x: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
    """EXPERIMENTAL halo-tile forward convolution (3x3, stride 1, pad 1; csrc/conv_halo.cu): compiled but
    not yet validated on hardware — nothing calls it by default."""
    assert x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16
    xv, wv = _nhwc(x), _nhwc(weight)
    n, h, w, cin = xv.shape
    cout, r, s, _ = wv.shape
    y = torch.empty((n, h, w, cout), dtype=torch.bfloat16, device=x.device)
    check(lib().tb200_conv2d_fprop_halo(xv.data_ptr(), wv.data_ptr(), y.data_ptr(), n, h, w, cin, cout, r, s,
                                        stream_ptr()))
    return y.permute(0, 3, 1, 2)


def dgrad_weight(weight: torch.Tensor) -> torch.Tensor:
    """The filter that turns a stride-1 convolution's input gradient into a forward convolution of dy:
    ``dx = conv(dy, dgrad_weight(w), padding = R-1-pad)`` — taps flipped, in / out channels swapped.
    ``weight`` [K,C,R,S] -> [C,K,R,S], channels_last."""
    return weight.flip(2, 3).transpose(0, 1).contiguous(memory_format=torch.channels_last)


def conv2d_dgrad_halo(dy: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
    """EXPERIMENTAL: input gradient of a 3x3 / stride 1 / pad 1 convolution on the halo-tile kernel (a
    forward pass over dy with the flipped, channel-swapped filter; the filter is rewritten per call, which
    is one small kernel — weights change every step anyway)."""
    return conv2d_fprop_halo(dy, dgrad_weight(weight))


class ConvStats:
    """Per-channel (sum, sum of squares) partials a convolution's epilogue wrote next to its output:
    ``partials`` fp32 [nparts, C, 2] over ``count`` values per channel. A following fused BatchNorm turns
    them into its scale / shift with one tiny kernel (``bn_finalize``) instead of re-reading the activation."""

    __slots__ = ("partials", "nparts", "count")

    def __init__(self, partials: torch.Tensor, nparts: int, count: int):
        self.partials, self.nparts, self.count = partials, nparts, count

    def sums(self) -> torch.Tensor:
        """[C, 2] totals in fp64 (tests / debugging)."""
        return self.partials[: self.nparts].double().sum(0)


def conv_stats_buffer(cout: int, device) -> torch.Tensor:
    parts = lib().tb200_conv_stats_floats(cout) // (2 * cout)
    return torch.empty((parts, cout, 2), dtype=torch.float32, device=device)


def conv3x3_fprop(x: torch.Tensor, weight: torch.Tensor, stats: bool = False):
    """3x3 / stride 1 / pad 1 convolution forward on the halo-tile kernel (tb200_conv3x3_fprop). ``x``
    [N,Cin,H,W] and ``weight`` [Cout,Cin,3,3] bf16 channels_last. With ``stats`` returns ``(y, ConvStats)``."""
    assert x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16
    xv, wv = _nhwc(x), _nhwc(weight)
    n, h, w, cin = xv.shape
    cout = wv.shape[0]
    assert wv.shape[1:] == (3, 3, cin)
    y = torch.empty((n, h, w, cout), dtype=torch.bfloat16, device=x.device)
    nparts = C.c_int(0)
    buf = conv_stats_buffer(cout, x.device) if stats else None
    check(lib().tb200_conv3x3_fprop(xv.data_ptr(), wv.data_ptr(), y.data_ptr(), n, h, w, cin, cout,
                                    buf.data_ptr() if stats else 0, buf.numel() if stats else 0,
                                    C.byref(nparts), stream_ptr()))
    y = y.permute(0, 3, 1, 2)
    return (y, ConvStats(buf[: nparts.value], nparts.value, n * h * w)) if stats else y


def conv3x3_supported(n: int, h: int, w: int, c_red: int, k_out: int, dgrad: bool = False) -> bool:
    """Do the halo-tile kernels serve this 3x3 / stride 1 / pad 1 geometry (no GPU work)?"""
    return lib().tb200_conv3x3_supported(n, h, w, c_red, k_out, int(dgrad)) == 1


def conv3x3_fprop_bias(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], relu: bool) -> torch.Tensor:
    """``relu?(conv3x3(x, weight) + bias)`` with bias and ReLU in the halo-tile kernel's epilogue."""
    assert x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16
    xv, wv = _nhwc(x), _nhwc(weight)
    n, h, w, cin = xv.shape
    cout = wv.shape[0]
    assert wv.shape[1:] == (3, 3, cin)
    y = torch.empty((n, h, w, cout), dtype=torch.bfloat16, device=x.device)
    bptr, bdt = (bias.data_ptr(), dtype_code(bias.dtype)) if bias is not None else (0, 0)
    check(lib().tb200_conv3x3_fprop_bias(xv.data_ptr(), wv.data_ptr(), bptr, bdt, int(relu), y.data_ptr(), n, h, w, cin,
                                         cout, stream_ptr()))
    return y.permute(0, 3, 1, 2)


def conv3x3_dgrad(dy: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
    """Input gradient of a 3x3 / stride 1 / pad 1 convolution (tb200_conv3x3_dgrad): the halo-tile kernel over
    dy with flipped taps, reading the forward ``weight`` [Cout,Cin,3,3] (channels_last) as stored."""
    assert dy.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16
    dyv, wv = _nhwc(dy), _nhwc(weight)
    n, h, w, cout = dyv.shape
    cin = wv.shape[3]
    assert wv.shape[:3] == (cout, 3, 3)
    dx = torch.empty((n, h, w, cin), dtype=torch.bfloat16, device=dy.device)
    check(lib().tb200_conv3x3_dgrad(dyv.data_ptr(), wv.data_ptr(), dx.data_ptr(), n, h, w, cin, cout, stream_ptr()))
    return dx.permute(0, 3, 1, 2)


_wgrad_ws: dict = {}


def release_stream_caches(device, stream: torch.cuda.Stream) -> None:
    """Drop every per-(device, stream) scratch this package keeps (BatchNorm / LayerNorm scratch, the 3x3 wgrad
    split planes, the native split-K workspace). Called by ``Trainer.close()``: a worker process runs many jobs, each
    on streams of its own, and the scratch of finished jobs must not accumulate. Nothing may be in flight."""
    dev = device.index if getattr(device, "index", None) is not None else torch.cuda.current_device()
    sp = stream.cuda_stream
    drop_python_stream_caches(dev, sp)
    with torch.cuda.device(dev):
        check(lib().tb200_gemm_release_stream(sp))


def drop_python_stream_caches(dev: int, stream_ptr_: int) -> int:
    """The Python half of :func:`release_stream_caches`: every cache is keyed ``(device index, stream handle, ...)``.
    Returns the number of entries dropped."""
    from . import batchnorm, layernorm
    dropped = 0
    for cache in (batchnorm._scratch, layernorm._scratch, _wgrad_ws):
        for key in [k for k in cache if k[0] == dev and k[1] == stream_ptr_]:
            del cache[key]
            dropped += 1
    return dropped


def conv3x3_wgrad_eligible(x: torch.Tensor, cout: int) -> bool:
    """Does the halo-tile weight-gradient kernel (tb200_conv3x3_wgrad) serve this 3x3 / stride 1 / pad 1 problem?"""
    n, cin, h, w = x.shape
    return lib().tb200_conv3x3_wgrad_workspace(n, h, w, cin, cout) > 0


def conv3x3_wgrad(x: torch.Tensor, dy: torch.Tensor) -> torch.Tensor:
    """Weight gradient of a 3x3 / stride 1 / pad 1 convolution on the halo-tile kernel (tb200_conv3x3_wgrad):
    x and dy are read once per half of the taps instead of once per tap / per output tile. Returns
    [Cout,Cin,3,3] channels_last. The fp32 split planes live in a scratch cached per (device, stream)."""
    assert x.dtype == torch.bfloat16 and dy.dtype == torch.bfloat16
    xv, dyv = _nhwc(x), _nhwc(dy)
    n, h, w, cin = xv.shape
    cout = dyv.shape[3]
    need = lib().tb200_conv3x3_wgrad_workspace(n, h, w, cin, cout)
    if need == 0:
        raise NativeError(f"conv3x3_wgrad: shape not served ({tuple(x.shape)} -> {cout}): " +
                          lib().tb200_last_error().decode(errors="replace"))
    key = (x.device.index, torch.cuda.current_stream(x.device).cuda_stream)
    ws = _wgrad_ws.get(key)
    if ws is None or ws.numel() * 4 < need:
        ws = torch.empty((need + 3) // 4, dtype=torch.float32, device=x.device)
        _wgrad_ws[key] = ws
    dw = torch.empty((cout, 3, 3, cin), dtype=torch.bfloat16, device=x.device)
    check(lib().tb200_conv3x3_wgrad(xv.data_ptr(), dyv.data_ptr(), dw.data_ptr(), n, h, w, cin, cout, ws.data_ptr(),
                                    ws.numel() * 4, stream_ptr()))
    return dw.permute(0, 3, 1, 2)


def conv2d_dgrad(dy: torch.Tensor, weight: torch.Tensor, x_shape, stride: int, pad: int) -> torch.Tensor:
    """Gradient w.r.t. the input of a stride-1 convolution (tb200_conv2d_dgrad)."""
    dyv, wv = _nhwc(dy), _nhwc(weight)
    n, cin, h, w = x_shape
    cout, r, s, _ = wv.shape
    dx = torch.empty((n, h, w, cin), dtype=torch.bfloat16, device=dy.device)
    check(lib().tb200_conv2d_dgrad(dyv.data_ptr(), wv.data_ptr(), dx.data_ptr(), n, h, w, cin, cout, r, s,
                                   stride, pad, stream_ptr()))
    return dx.permute(0, 3, 1, 2)


def conv2d_wgrad(x: torch.Tensor, dy: torch.Tensor, kernel, stride: int, pad: int,
                 out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """Gradient w.r.t. the weight (tb200_conv2d_wgrad); returns [Cout,Cin,R,S] channels_last."""
    xv, dyv = _nhwc(x), _nhwc(dy)
    n, h, w, cin = xv.shape
    cout = dyv.shape[3]
    r, s = kernel
    dw = torch.empty((cout, r, s, cin), dtype=out_dtype, device=x.device)
    check(lib().tb200_conv2d_wgrad(xv.data_ptr(), dyv.data_ptr(), dw.data_ptr(), n, h, w, cin, cout, r, s,
                                   stride, pad, dtype_code(out_dtype), stream_ptr()))
    return dw.permute(0, 3, 1, 2)


def transpose(x: torch.Tensor) -> torch.Tensor:
    """bf16 [R,C] -> [C,R] through the tiled transpose kernel."""
    assert x.dtype == torch.bfloat16 and x.is_contiguous() and x.dim() == 2
    r, c = x.shape
    out = torch.empty((c, r), dtype=x.dtype, device=x.device)
    check(lib().tb200_transpose_bf16(x.data_ptr(), out.data_ptr(), r, c, stream_ptr()))
    return out


def u8_to_bf16_norm(src: torch.Tensor, mean: Sequence[float], std: Sequence[float],
                    out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """uint8 NHWC batch -> normalised bf16 NHWC."""
    assert src.dtype == torch.uint8 and src.is_contiguous() and src.shape[-1] == 3
    if out is None:
        out = torch.empty(src.shape, dtype=torch.bfloat16, device=src.device)
    m = (C.c_float * 3)(*mean)
    s = (C.c_float * 3)(*[1.0 / v for v in std])
    check(lib().tb200_u8_to_bf16_norm(src.data_ptr(), out.data_ptr(), src.numel() // 3, m, s,
                                      stream_ptr()))
    return out


def u8_to_bf16_norm_pad8(src: torch.Tensor, mean: Sequence[float], std: Sequence[float],
                         out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """uint8 NHWC batch [..., 3] -> normalised bf16 NHWC [..., 8] (channels 3..7 zero)."""
    assert src.dtype == torch.uint8 and src.is_contiguous() and src.shape[-1] == 3
    if out is None:
        out = torch.empty((*src.shape[:-1], 8), dtype=torch.bfloat16, device=src.device)
    m = (C.c_float * 3)(*mean)
    s = (C.c_float * 3)(*[1.0 / v for v in std])
    check(lib().tb200_u8_to_bf16_norm_pad8(src.data_ptr(), out.data_ptr(), src.numel() // 3, m, s,
                                           stream_ptr()))
    return out


def stem_packed_shape(n: int, h: int, w: int):
    """Shape [n, oh+3, ow, 64] of the packed stem input for an [n, h, w, 3] image batch (csrc/stem.cu)."""
    return (n, (h - 1) // 2 + 1 + 3, (w - 1) // 2 + 1, 64)


def stem7x7_pack(src: torch.Tensor, mean: Sequence[float], std: Sequence[float],
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """uint8 NHWC batch -> the packed, normalised bf16 tensor Y [n, oh+3, ow, 64] over which the 7x7 / stride 2
    stem convolution is a 4x1 / stride 1 convolution with 64 input channels (tb200_stem7x7_pack)."""
    assert src.dtype == torch.uint8 and src.is_contiguous() and src.dim() == 4 and src.shape[-1] == 3
    n, h, w, _ = src.shape
    if out is None:
        out = torch.empty(stem_packed_shape(n, h, w), dtype=torch.bfloat16, device=src.device)
    assert tuple(out.shape) == stem_packed_shape(n, h, w) and out.is_contiguous()
    m = (C.c_float * 3)(*mean)
    s = (C.c_float * 3)(*[1.0 / v for v in std])
    check(lib().tb200_stem7x7_pack(src.data_ptr(), out.data_ptr(), n, h, w, m, s, stream_ptr()))
    return out


def stem7x7_weight_pack(weight: torch.Tensor) -> torch.Tensor:
    """[cout,3,7,7] bf16 channels_last stem weight -> the [cout,64,4,1] channels_last weight of the packed
    convolution (tb200_stem7x7_weight_pack)."""
    wv = _nhwc(weight)
    cout = wv.shape[0]
    assert wv.shape[1:] == (7, 7, 3) and weight.dtype == torch.bfloat16
    wp = torch.empty((cout, 4, 1, 64), dtype=torch.bfloat16, device=weight.device)
    check(lib().tb200_stem7x7_weight_pack(wv.data_ptr(), wp.data_ptr(), cout, stream_ptr()))
    return wp.permute(0, 3, 1, 2)


def stem7x7_wgrad_unpack(dwp: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """fp32 gradient of the packed weight [cout,64,4,1] (channels_last) -> [cout,3,7,7] channels_last."""
    v = _nhwc(dwp)
    cout = v.shape[0]
    assert v.shape[1:] == (4, 1, 64) and dwp.dtype == torch.float32
    dw = torch.empty((cout, 7, 7, 3), dtype=dtype, device=dwp.device)
    check(lib().tb200_stem7x7_wgrad_unpack(v.data_ptr(), dw.data_ptr(), cout, dtype_code(dtype), stream_ptr()))
    return dw.permute(0, 3, 1, 2)


def sumsq(x: torch.Tensor, scratch: Optional[torch.Tensor] = None,
          out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Deterministic sum of squares (gradient-norm watchdog)."""
    nfl = int(lib().tb200_sumsq_scratch_floats())
    if scratch is None:
        scratch = torch.empty(nfl, dtype=torch.float32, device=x.device)
    if out is None:
        out = torch.empty(1, dtype=torch.float32, device=x.device)
    assert scratch.numel() >= nfl
    check(lib().tb200_sumsq(x.data_ptr(), dtype_code(x.dtype), x.numel(), scratch.data_ptr(),
                            out.data_ptr(), stream_ptr()))
    return out
