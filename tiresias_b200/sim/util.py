"""Small helpers with the reference's names (core/util.py:12-56, C14).

``print_fn`` funnels into :mod:`logging` (INFO / WARNING / ERROR-and-exit); the simulator's own
logger is ``tiresias_b200.sim`` so several runs in one process can attach their own handlers.
"""
from __future__ import annotations

import logging
import os
import sys
from typing import Any, Iterable, Mapping, Optional

LOG_LEVEL_INFO = 0
LOG_LEVEL_WARNING = 1
LOG_LEVEL_ERROR = 2

logger = logging.getLogger("tiresias_b200.sim")


def print_fn(log: Any, level: int = LOG_LEVEL_INFO) -> None:
    if level == LOG_LEVEL_INFO:
        logger.info(log)
    elif level == LOG_LEVEL_WARNING:
        logger.warning(log)
    elif level == LOG_LEVEL_ERROR:
        logger.error(log)
        sys.exit(1)


_UNITS = {"GiB": 3, "MiB": 2, "KiB": 1}


def convert_bytes(num_byte, unit: str = "GiB"):
    """Successive divisions by 1024 (kept as divisions so results match the reference bit for
    bit; core/util.py:21-29)."""
    for _ in range(_UNITS.get(unit, 0)):
        num_byte = num_byte / 1024
    return num_byte


def make_dir_if_not_exist(local_path: str) -> None:
    if not os.path.exists(local_path):
        os.makedirs(local_path)
        logger.info("Created directory %s", local_path)


def mkdir(folder_path: str) -> None:
    os.makedirs(folder_path, exist_ok=True)


def search_dict_list(dict_list: Iterable[Mapping], key, value) -> Optional[Mapping]:
    """First mapping in ``dict_list`` with ``entry[key] == value`` (core/util.py:44-56)."""
    for entry in dict_list:
        if key in entry and entry[key] == value:
            return entry
    return None
