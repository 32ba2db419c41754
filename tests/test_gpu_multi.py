"""Multi-GPU checks of the in-kernel communication paths over real NVLink peer memory.

Launches tests/multi_worker.py with one process per GPU (torchrun, NCCL bootstrap on 127.0.0.1)
for every world size in {2, 4, 8} that the box has GPUs for. Skipped on single-GPU boxes; the
single-GPU suite covers the same kernels with emulated ranks (test_gpu_kernels.py).
"""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 4, 8])
def test_in_kernel_collectives(native_lib, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, box has {torch.cuda.device_count()}")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "multi_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=420, cwd=ROOT)
    tail = (res.stdout + "\n" + res.stderr)[-4000:]
    assert res.returncode == 0, tail
    assert "MULTI OK" in res.stdout, tail


def test_nvls_multimem_primitives(native_lib):
    """multimem.ld_reduce / multimem.st through an NVLS multicast object: sums and broadcasts must be exact
    (small integers in bf16). Single process, two GPUs (the multi-process path over 2/4/8 GPUs is covered by
    multi_worker.py through the trainers)."""
    from tiresias_b200.ops import native
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ndev = 2
    if not native.nvls_supported(0):
        pytest.skip("this box has no NVLS multicast support")
    rep = native.nvls_selftest(ndev, 32 << 20)
    if rep is None:
        pytest.skip("NVLS unavailable: " + native.lib().tb200_last_error().decode(errors="replace"))
    assert rep["ok"] and "error" not in rep, rep
    want = ndev * (ndev + 1) / 2
    assert rep["reduce_first"] == want and rep["bcast_min"] == want == rep["bcast_max"]
    assert rep["allreduce_min"] == want == rep["allreduce_max"]
