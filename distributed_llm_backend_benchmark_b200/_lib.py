"""ctypes binding of libdlbb.so (the C ABI declared in include/dlbb.h and include/dlbb_comm.h).

The library is built in-tree by ``build.py`` (``python build.py`` or ``__graft_entry__.build()``)
so the ``.so`` travels with the source snapshot to the GPU box.  On a machine with a GPU a
missing library is a hard error: ops never fall back silently to PyTorch there.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "lib" / "libdlbb.so"

_lib = None
_load_error: str | None = None

c_void_p, c_int, c_int64, c_float, c_size_t = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t
c_uint32, c_uint64 = C.c_uint32, C.c_uint64

_SIGNATURES = {
    "dlbb_last_error": (C.c_char_p, []),
    "dlbb_launch_count": (c_uint64, []),
    "dlbb_build_info": (C.c_char_p, []),
    "dlbb_gemm_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_int64, c_int64,
                               c_int, c_int, c_int, c_void_p, c_int64, c_void_p]),
    "dlbb_gemm_bf16_simt": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_int64,
                                    c_int64, c_int, c_int, c_int, c_void_p, c_int64, c_void_p]),
    "dlbb_gemm_set_tile_n": (None, [c_int]),
    "dlbb_gemm_set_pair_mode": (None, [c_int]),
    "dlbb_layernorm_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                   c_float, c_void_p]),
    "dlbb_layernorm_bwd_scratch_bytes": (c_size_t, [c_int, c_int]),
    "dlbb_layernorm_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "dlbb_layernorm_bwd_ex": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "dlbb_gelu_fwd": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "dlbb_gelu_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "dlbb_add": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "dlbb_copy2d": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_void_p]),
    "dlbb_cast": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int64, c_void_p]),
    "dlbb_mse_scratch_bytes": (c_size_t, [c_int64]),
    "dlbb_mse_loss": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_void_p]),
    "dlbb_adam_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_float,
                               c_float, c_float, c_float, c_float, c_int, c_void_p]),
    "dlbb_adam_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_float, c_float,
                              c_float, c_float, c_float, c_int, c_void_p]),
    "dlbb_fill_u32": (c_int, [c_void_p, c_uint32, c_int64, c_void_p]),
    "dlbb_adam_set_max_ctas": (None, [c_int]),
}

# Optional symbols (collectives / bootstrap) are bound lazily by parallel/symm.py through bind().


class DlbbError(RuntimeError):
    pass


def bind(lib, name: str, restype, argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = argtypes
    return fn


def load(required: bool = True):
    """Return the loaded library (cached).  ``required=False`` returns None when unavailable."""
    global _lib, _load_error
    if _lib is not None:
        return _lib
    if _load_error is not None and not required:
        return None
    path = Path(os.environ.get("DLBB_LIB", str(LIB_PATH)))
    try:
        if not path.exists():
            raise OSError(f"{path} not found - run `python build.py` first")
        lib = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
        for name, (res, args) in _SIGNATURES.items():
            bind(lib, name, res, args)
        _lib = lib
        return lib
    except OSError as e:  # missing file, missing libcudart, ...
        _load_error = str(e)
        if required:
            raise DlbbError(f"cannot load libdlbb.so: {e}") from e
        return None


def available() -> bool:
    return load(required=False) is not None


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().dlbb_last_error().decode(errors="replace")
        raise DlbbError(f"{what or 'dlbb call'} failed ({rc}): {msg}")


def launch_count() -> int:
    lib = load(required=False)
    return int(lib.dlbb_launch_count()) if lib is not None else 0
