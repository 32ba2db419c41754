"""Worker of tests/test_exec.py::test_fd_exchange: every rank offers a pipe whose content names it and
checks that it received every peer's pipe (file descriptors passed as SCM_RIGHTS over unix sockets)."""
import os
import sys

import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tiresias_b200.parallel.symm import exchange_fds  # noqa: E402


def main() -> None:
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    pipes = []
    for peer in range(world):                 # one pipe per reader, so that nobody steals a message
        r, w = os.pipe()
        os.write(w, f"from {rank} to {peer}".encode())
        pipes.append((r, w))
    got = exchange_fds(rank, world, [r for r, _ in pipes])
    for peer, fds in sorted(got.items()):
        assert len(fds) == world, (peer, fds)
        assert os.read(fds[rank], 100).decode() == f"from {peer} to {rank}"
        for f in fds:
            os.close(f)
    dist.barrier()
    if rank == 0:
        print("FD OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
