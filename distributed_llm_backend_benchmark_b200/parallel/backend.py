"""Registration of the native process-group backend with torch.distributed.

``register()`` makes ``dist.init_process_group(backend="dlbb")`` create the C++ ``BackendDlbb``
(csrc/pg/backend.cpp) for both CUDA and CPU tensors.  ``register(alias)`` adds further names for the
same backend - the ``oneccl_bindings_for_pytorch`` shim registers ``"ccl"`` this way so the
reference's ``init_process_group(backend='ccl')`` (test/ccl.py:24, test/test_deepseed.py:24,
collectives/1d/dsccl.py:16) needs no edit.

On a machine without a GPU the same names resolve to Gloo, which is all the host-side plumbing
tests need (BASELINE.json config 1).
"""
from __future__ import annotations

import importlib.util
import os
from pathlib import Path

import torch
import torch.distributed as dist

_EXT = None
_registered: set[str] = set()
NATIVE_NAME = "dlbb"


def load_extension():
    global _EXT
    if _EXT is None:
        path = Path(__file__).resolve().parent.parent / "lib" / "_dlbb_pg.so"
        if not path.exists():
            raise RuntimeError(f"{path} not found - run `python build.py`")
        spec = importlib.util.spec_from_file_location("_dlbb_pg", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _EXT = mod
    return _EXT


def use_native() -> bool:
    return torch.cuda.is_available() and os.environ.get("DLBB_FORCE_GLOO", "0") != "1"


def _create(store, rank, size, timeout):
    if use_native():
        local = int(os.environ.get("LOCAL_RANK", rank % max(torch.cuda.device_count(), 1)))
        if torch.cuda.current_device() != local and "DLBB_KEEP_DEVICE" not in os.environ:
            torch.cuda.set_device(local)
        return load_extension().create_backend(store, rank, size, timeout)
    # CPU-only host: plumbing over Gloo
    return dist.ProcessGroupGloo(store, rank, size, timeout=timeout)


def register(name: str = NATIVE_NAME) -> str:
    """Idempotently register `name` as a torch.distributed backend served by this framework."""
    if name not in _registered:
        try:
            dist.Backend.register_backend(name, _create, devices=["cpu", "cuda"])
        except (RuntimeError, ValueError) as e:  # name taken (e.g. a real oneCCL install): leave it alone
            if "already" not in str(e).lower():
                raise
        _registered.add(name)
    return name


def native_backend(group=None):
    """The BackendDlbb object behind a process group (for knobs such as set_ctas / set_algo)."""
    group = group or dist.group.WORLD
    dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    return group._get_backend(dev)


def init_process_group(backend: str | None = None, **kw):
    """``dist.init_process_group`` with rendezvous defaults for a single node (127.0.0.1)."""
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    if backend in (None, NATIVE_NAME, "ccl"):
        backend = register(backend or NATIVE_NAME)
    if not dist.is_initialized():
        dist.init_process_group(backend=backend, **kw)
    return dist.group.WORLD
