"""Tracing / logging helpers.

The reference only has host wall-clock timers and backend log switches in the environment
(``CCL_LOG_LEVEL=debug``, ``CCL_LOG_FILE=ccl_log_rank_%r.txt``, ``I_MPI_DEBUG`` - collectives/3d/launch_dsccl.sh:34,
collectives/3d/launch_mpiccl.sh:12,17-18).  Analogues here:

* ``DLBB_LOG_LEVEL=info|debug`` - the communicator, launcher and engines print what they set up;
  ``DLBB_LOG_FILE=dlbb_log_rank_%r.txt`` redirects it per rank (``%r`` = rank);
* ``DLBB_NVTX=1`` - NVTX ranges around every layer / bucket / collective (``nvtx_range``), visible in Nsight Systems;
* device timing is done with CUDA events (``utils.metrics.DeviceTimer``) and reduced as max over ranks.
"""
from __future__ import annotations

import contextlib
import os
import sys

_LEVELS = {"": 0, "error": 0, "warn": 1, "warning": 1, "info": 2, "debug": 3}
_level = _LEVELS.get(os.environ.get("DLBB_LOG_LEVEL", "").lower(), 2 if os.environ.get("DLBB_LOG_LEVEL") else 0)
_nvtx = os.environ.get("DLBB_NVTX", "0") == "1"
_stream = None


def _out():
    global _stream
    if _stream is None:
        path = os.environ.get("DLBB_LOG_FILE")
        if path:
            _stream = open(path.replace("%r", os.environ.get("RANK", "0")), "a", buffering=1)
        else:
            _stream = sys.stderr
    return _stream


def log(level: str, msg: str) -> None:
    if _LEVELS.get(level, 2) <= _level:
        print(f"[dlbb r{os.environ.get('RANK', '0')} {level}] {msg}", file=_out())


def info(msg: str) -> None:
    log("info", msg)


def debug(msg: str) -> None:
    log("debug", msg)


@contextlib.contextmanager
def nvtx_range(name: str):
    """NVTX range when DLBB_NVTX=1 and CUDA is present; free otherwise."""
    if _nvtx:
        import torch
        if torch.cuda.is_available():
            torch.cuda.nvtx.range_push(name)
            try:
                yield
            finally:
                torch.cuda.nvtx.range_pop()
            return
    yield
