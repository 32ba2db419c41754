"""GPU placement inside one NVSwitch domain.

Behind NVSwitch every GPU pair has the same bandwidth, so the reference's rack/node distance
(infra/infrastructure.py:135-147) collapses; what still matters is that a job's ranks form a group the
worker pool has a communicator for. The pool pre-builds one group per *aligned power-of-two block*
(8 singles, 4 pairs, 2 quads, 1 octet on 8 GPUs), and this allocator hands out exactly those blocks:
a buddy allocator over GPU indices.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

Block = Tuple[int, int]          # (first GPU, number of GPUs)


def normalize_gpus(num_gpu: int, total: int) -> int:
    """Largest power of two <= min(num_gpu, total): a trace job asking for 12 or 32 GPUs is executed
    on 8 of an 8-GPU box (the down-scaled spec of SURVEY §7.4 item 9)."""
    g = max(1, min(int(num_gpu), total))
    p = 1
    while p * 2 <= g:
        p *= 2
    return p


def all_blocks(total: int) -> List[Block]:
    out, size = [], 1
    while size <= total:
        out += [(s, size) for s in range(0, total, size)]
        size *= 2
    return out


class BuddyAllocator:
    def __init__(self, total: int = 8):
        if total & (total - 1):
            raise ValueError("GPU count must be a power of two")
        self.total = total
        self.busy = [False] * total
        self.owner: Dict[Block, str] = {}

    def free_gpus(self) -> int:
        return self.busy.count(False)

    def can_alloc(self, size: int) -> bool:
        return self._find(size) is not None

    def _find(self, size: int) -> Optional[Block]:
        """Best fit: the free aligned block of ``size`` whose enclosing 2*size block is the fullest, so
        large aligned blocks stay free for large jobs."""
        best, best_key = None, None
        for start in range(0, self.total, size):
            if any(self.busy[start:start + size]):
                continue
            parent = (start // (2 * size)) * 2 * size
            crowd = sum(self.busy[parent:parent + 2 * size]) if 2 * size <= self.total else 0
            key = (-crowd, start)
            if best_key is None or key < best_key:
                best, best_key = (start, size), key
        return best

    def alloc(self, size: int, owner: str = "") -> Optional[Block]:
        blk = self._find(size)
        if blk is None:
            return None
        for g in range(blk[0], blk[0] + blk[1]):
            self.busy[g] = True
        self.owner[blk] = owner
        return blk

    def free(self, blk: Block) -> None:
        for g in range(blk[0], blk[0] + blk[1]):
            self.busy[g] = False
        self.owner.pop(blk, None)
