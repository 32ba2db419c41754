#!/usr/bin/env python
"""Hardware probe for halo-tile convolutions (csrc/gemm_tcgen05.cu probe_shift_kernel): does a K-major,
128-byte-swizzled shared-memory operand still read correctly when its descriptor starts `shift` rows into
the tile, and which "matrix base offset" makes it so? Prints one JSON line per (shift, base_offset)."""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tiresias_b200.ops import native  # noqa: E402


def main():
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    lib = native.lib()
    lib.tb200_probe_shifted_operand.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    lib.tb200_probe_shifted_operand.restype = C.c_int
    torch.manual_seed(0)
    a = torch.randn(256, 64, device=dev).to(torch.bfloat16)
    b = torch.randn(64, 64, device=dev).to(torch.bfloat16)
    full = a.float() @ b.float().t()                       # [256, 64]: row j = what source row j produces
    out = torch.empty(128, 64, device=dev)
    rows = []
    for shift in (0, 1, 2, 3, 7, 8, 9, 16, 30, 58, 59, 116):
        for bo in sorted({0, shift & 7}):
            out.zero_()
            native.check(lib.tb200_probe_shifted_operand(a.data_ptr(), b.data_ptr(), out.data_ptr(), shift, bo,
                                                         native.stream_ptr()))
            torch.cuda.synchronize()
            want = full[shift:shift + 128]
            err = float((out - want).abs().max())
            # which source row does each output row look like?
            d = torch.cdist(out, full)                     # [128, 256]
            src = d.argmin(dim=1)
            exact = int((d.min(dim=1).values < 1e-2).sum())
            rows.append({"shift": shift, "base_offset": bo, "max_abs_err": round(err, 4), "ok": err < 0.05,
                         "rows_matching_some_source_row": exact, "first_sources": src[:12].tolist()})
            print(json.dumps(rows[-1]), flush=True)
    # MN-major operand shifted along the reduction dimension (halo-tile wgrad)
    if hasattr(lib, "tb200_probe_shifted_mn_operand"):
        lib.tb200_probe_shifted_mn_operand.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        lib.tb200_probe_shifted_mn_operand.restype = C.c_int
        a2 = torch.randn(128, 64, device=dev).to(torch.bfloat16)
        b2 = torch.randn(256, 64, device=dev).to(torch.bfloat16)
        for shift in (0, 1, 3, 8, 9, 18, 66, 130, 192):
            out.zero_()
            native.check(lib.tb200_probe_shifted_mn_operand(a2.data_ptr(), b2.data_ptr(), out.data_ptr(), shift,
                                                            native.stream_ptr()))
            torch.cuda.synchronize()
            want = a2.float() @ b2[shift:shift + 64].float()
            err = float((out - want).abs().max())
            rows.append({"operand": "B MN-major", "shift": shift, "max_abs_err": round(err, 4), "ok": err < 0.05})
            print(json.dumps(rows[-1]), flush=True)
    # TMA boxes landing at 128-byte (not 1024-byte) aligned shared-memory offsets: needed for halo tiles whose
    # padded rows are not multiples of 8 pixels
    if hasattr(lib, "tb200_probe_tma_dst_shift"):
        lib.tb200_probe_tma_dst_shift.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        lib.tb200_probe_tma_dst_shift.restype = C.c_int
        a3 = torch.randn(128, 64, device=dev).to(torch.bfloat16)
        want = a3.float() @ b.float().t()
        for shift in (0, 1, 2, 5, 8, 9, 30, 58, 120):
            for piece in (128, 2, 1):
                out.zero_()
                native.check(lib.tb200_probe_tma_dst_shift(a3.data_ptr(), b.data_ptr(), out.data_ptr(), shift, piece,
                                                           native.stream_ptr()))
                torch.cuda.synchronize()
                err = float((out - want).abs().max())
                rows.append({"probe": "tma_dst", "shift": shift, "piece_rows": piece, "max_abs_err": round(err, 4),
                             "ok": err < 0.05})
                print(json.dumps(rows[-1]), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "probe_shift.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
