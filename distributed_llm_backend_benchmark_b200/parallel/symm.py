"""Python face of the native communicator (include/dlbb_comm.h): symmetric memory, collectives on
torch tensors, and the fused data-parallel optimizer step.

One ``SymmComm`` per process (one process per GPU).  All ranks must create communicators, allocate
symmetric tensors and issue collectives in the same order - the usual SPMD contract of
torch.distributed / MPI, which this replaces for the GPU path (ref: deepspeed.comm call sites in
collectives/1d/dsccl.py:59-163 and mpi4py call sites in collectives/1d/openmpi.py:55-198).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from .. import _lib

VP, I, I64, F, SZ = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t

DTYPES = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2, torch.float64: 3, torch.uint8: 4, torch.int8: 4,
          torch.int32: 5, torch.int64: 6}
OPS = {"sum": 0, "max": 1, "min": 2, "prod": 3}
ALGOS = {"auto": 0, "oneshot": 1, "twoshot": 2, "nvls": 3, "direct": 4}


class _View(C.Structure):
    _fields_ = [("ndim", I), ("size", I64 * 4), ("stride", I64 * 4)]


_SIGS = {
    "dlbb_comm_create": (I, [I, I, I, C.c_char_p, SZ, C.POINTER(VP)]),
    "dlbb_comm_destroy": (I, [VP]),
    "dlbb_comm_has_multicast": (I, [VP]),
    "dlbb_comm_host_barrier": (I, [VP]),
    "dlbb_comm_host_allgather": (I, [VP, VP, VP, SZ]),
    "dlbb_symm_alloc": (I, [VP, SZ, C.POINTER(I)]),
    "dlbb_symm_ptr": (VP, [VP, I, I]),
    "dlbb_symm_mc_ptr": (VP, [VP, I]),
    "dlbb_symm_size": (SZ, [VP, I]),
    "dlbb_symm_free": (I, [VP, I]),
    "dlbb_comm_set_ctas": (None, [VP, I]),
    "dlbb_comm_set_oneshot_max_bytes": (None, [VP, SZ]),
    "dlbb_comm_set_zero_copy_min_bytes": (None, [VP, SZ]),
    "dlbb_comm_zero_copy_min_bytes": (SZ, [VP]),
    "dlbb_barrier": (I, [VP, VP]),
    "dlbb_allreduce": (I, [VP, VP, VP, I64, I, I, I, C.POINTER(_View), VP]),
    "dlbb_allreduce_v": (I, [VP, VP, VP, I64, I, I, I, C.POINTER(_View), C.POINTER(_View), VP]),
    "dlbb_allreduce_multi": (I, [VP, C.POINTER(VP), C.POINTER(I64), I, I, I, VP]),
    "dlbb_broadcast_v": (I, [VP, VP, I64, I, I, C.POINTER(_View), VP]),
    "dlbb_reduce_v": (I, [VP, VP, VP, I64, I, I, I, C.POINTER(_View), C.POINTER(_View), VP]),
    "dlbb_allreduce_residual_bf16": (I, [VP, VP, VP, VP, I64, VP]),
    "dlbb_reduce_scatter": (I, [VP, VP, VP, I64, I, I, VP]),
    "dlbb_allgather": (I, [VP, VP, VP, I64, I, C.POINTER(_View), VP]),
    "dlbb_broadcast": (I, [VP, VP, I64, I, I, VP]),
    "dlbb_reduce": (I, [VP, VP, VP, I64, I, I, I, VP]),
    "dlbb_gather": (I, [VP, VP, VP, I64, I, I, VP]),
    "dlbb_scatter": (I, [VP, VP, VP, I64, I, I, VP]),
    "dlbb_alltoall": (I, [VP, VP, VP, I64, I, VP]),
    "dlbb_sendrecv": (I, [VP, VP, I64, I, VP, I64, I, I, VP]),
    "dlbb_fused_reduce_adam_bf16": (I, [VP, I, I, I64, I64, VP, VP, VP, F, F, F, F, F, F, I, I, VP]),
    "dlbb_gemm_allreduce_bf16": (I, [VP, VP, VP, I, I, I, I64, I64, I, I, I, I64, I, I64, VP, I64, I, VP]),
}
_bound = False


def _api():
    global _bound
    lib = _lib.load()
    if not _bound:
        for name, (res, args) in _SIGS.items():
            _lib.bind(lib, name, res, args)
        _bound = True
    return lib


class _RawCuda:
    """Minimal __cuda_array_interface__ carrier so torch can view VMM memory without copying."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


def _stream(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _strided_view(t: torch.Tensor):
    """(contiguous?, _View) for a tensor of up to 4 dims; dense tensors need no descriptor."""
    if t.is_contiguous():
        return None
    if t.dim() > 4:
        raise ValueError("in-kernel packing supports views of up to 4 dimensions")
    v = _View()
    v.ndim = t.dim()
    for i in range(t.dim()):
        v.size[i] = t.shape[i]
        v.stride[i] = t.stride(i)
    return v


class SymmComm:
    _seq = 0

    def __init__(self, rank: int, world_size: int, device, job: str, staging_mb: int | None = None):
        self.rank, self.world_size = rank, world_size
        self.device = torch.device(device)
        self.lib = _api()
        h = VP()
        stage = (staging_mb << 20) if staging_mb else 0
        torch.cuda.set_device(self.device)
        _lib.check(self.lib.dlbb_comm_create(rank, world_size, self.device.index, job.encode(), stage, C.byref(h)),
                   "comm_create")
        self.h = h
        self._segs: dict[int, torch.Tensor] = {}
        self._dp = None

    @classmethod
    def from_env(cls, device=None, job: str | None = None, **kw) -> "SymmComm":
        """Rank / world size from the launcher's environment (RANK, WORLD_SIZE, LOCAL_RANK, MASTER_PORT)."""
        rank = int(os.environ.get("RANK", "0"))
        world = int(os.environ.get("WORLD_SIZE", "1"))
        if device is None:
            device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank % max(torch.cuda.device_count(), 1))))
        if job is None:
            cls._seq += 1
            job = f"{os.environ.get('DLBB_JOB_ID', os.environ.get('MASTER_PORT', '29500'))}-{cls._seq}"
        return cls(rank, world, device, job, **kw)

    from_torch_distributed = from_env

    # ------------------------------------------------------------------ life cycle
    def destroy(self) -> None:
        if self.h:
            torch.cuda.synchronize(self.device)
            self._segs.clear()
            self.lib.dlbb_comm_destroy(self.h)
            self.h = None

    @property
    def has_multicast(self) -> bool:
        return bool(self.lib.dlbb_comm_has_multicast(self.h))

    def host_barrier(self) -> None:
        _lib.check(self.lib.dlbb_comm_host_barrier(self.h), "host_barrier")

    def set_ctas(self, n: int) -> None:
        self.lib.dlbb_comm_set_ctas(self.h, n)

    def set_oneshot_max_bytes(self, n: int) -> None:
        self.lib.dlbb_comm_set_oneshot_max_bytes(self.h, n)

    def set_zero_copy_min_bytes(self, n: int) -> None:
        """Messages of at least `n` bytes on plain (cudaMalloc / caching-allocator) tensors are exchanged directly
        between the ranks' own tensors instead of through the staging area (0 = never).  Must be set to the same
        value on every rank; such a call blocks the host until every rank has issued it."""
        self.lib.dlbb_comm_set_zero_copy_min_bytes(self.h, n)

    @property
    def zero_copy_min_bytes(self) -> int:
        """0 when zero-copy is unavailable (single rank, IPC self-test failed) or switched off."""
        return int(self.lib.dlbb_comm_zero_copy_min_bytes(self.h))

    # ------------------------------------------------------------------ symmetric tensors
    def alloc(self, numel: int, dtype: torch.dtype = torch.bfloat16) -> tuple[torch.Tensor, int]:
        """Collective.  Returns (tensor viewing this rank's copy, segment id)."""
        es = torch.empty((), dtype=dtype).element_size()
        seg = I()
        _lib.check(self.lib.dlbb_symm_alloc(self.h, numel * es, C.byref(seg)), "symm_alloc")
        ptr = self.lib.dlbb_symm_ptr(self.h, seg.value, self.rank)
        raw = torch.as_tensor(_RawCuda(ptr, numel * es), device=self.device)
        t = raw.view(dtype)
        self._segs[seg.value] = t
        return t, seg.value

    def peer_view(self, seg: int, peer: int, numel: int, dtype: torch.dtype) -> torch.Tensor:
        """A tensor over `peer`'s copy of a segment, mapped into this process (debugging / tests)."""
        es = torch.empty((), dtype=dtype).element_size()
        ptr = self.lib.dlbb_symm_ptr(self.h, seg, peer)
        return torch.as_tensor(_RawCuda(ptr, numel * es), device=self.device).view(dtype)

    # ------------------------------------------------------------------ collectives on torch tensors
    def barrier(self) -> None:
        _lib.check(self.lib.dlbb_barrier(self.h, _stream(self.device)), "barrier")

    def all_reduce(self, t: torch.Tensor, op: str = "sum", algo: str = "auto", out: torch.Tensor | None = None):
        """In place by default.  A non-contiguous input (<= 4 dims) is packed inside the kernel; the result is
        then written to `out` (contiguous), which is required in that case."""
        view = _strided_view(t)
        if out is None and view is not None:
            # in place on the strided view: gathered by the kernel's pack phase, scattered back by its unpack phase
            _lib.check(self.lib.dlbb_allreduce_v(self.h, t.data_ptr(), t.data_ptr(), t.numel(), DTYPES[t.dtype], OPS[op],
                                                 ALGOS[algo], C.byref(view), C.byref(view), _stream(self.device)),
                       "allreduce (strided)")
            return t
        if out is None:
            out = t
        assert out.is_contiguous() and out.numel() == t.numel()
        _lib.check(self.lib.dlbb_allreduce(self.h, t.data_ptr(), out.data_ptr(), t.numel(), DTYPES[t.dtype], OPS[op],
                                           ALGOS[algo], C.byref(view) if view is not None else None,
                                           _stream(self.device)), "allreduce")
        return out

    def all_reduce_residual(self, partial: torch.Tensor, residual: torch.Tensor, out: torch.Tensor):
        assert partial.dtype == torch.bfloat16 and partial.is_contiguous() and residual.is_contiguous()
        _lib.check(self.lib.dlbb_allreduce_residual_bf16(self.h, partial.data_ptr(), residual.data_ptr(),
                                                         out.data_ptr(), partial.numel(), _stream(self.device)),
                   "allreduce_residual")
        return out

    def reduce_scatter(self, inp: torch.Tensor, out: torch.Tensor, op: str = "sum"):
        assert inp.is_contiguous() and out.is_contiguous() and inp.numel() == out.numel() * self.world_size
        _lib.check(self.lib.dlbb_reduce_scatter(self.h, inp.data_ptr(), out.data_ptr(), out.numel(), DTYPES[inp.dtype],
                                                OPS[op], _stream(self.device)), "reduce_scatter")
        return out

    def all_gather(self, inp: torch.Tensor, out: torch.Tensor):
        view = _strided_view(inp)
        assert out.is_contiguous() and out.numel() == inp.numel() * self.world_size
        _lib.check(self.lib.dlbb_allgather(self.h, inp.data_ptr(), out.data_ptr(), inp.numel(), DTYPES[inp.dtype],
                                           C.byref(view) if view is not None else None, _stream(self.device)),
                   "allgather")
        return out

    def all_reduce_multi(self, tensors: list[torch.Tensor], op: str = "sum") -> None:
        """Bucket fusion: all-reduce many (small) contiguous tensors of one dtype in place with one launch per bucket."""
        if not tensors:
            return
        dt = tensors[0].dtype
        assert all(t.dtype == dt and t.is_contiguous() for t in tensors)
        ptrs = (VP * len(tensors))(*[t.data_ptr() for t in tensors])
        counts = (I64 * len(tensors))(*[t.numel() for t in tensors])
        _lib.check(self.lib.dlbb_allreduce_multi(self.h, ptrs, counts, len(tensors), DTYPES[dt], OPS[op],
                                                 _stream(self.device)), "allreduce_multi")

    def broadcast(self, t: torch.Tensor, root: int = 0):
        view = _strided_view(t)
        if view is not None:   # the root gathers the view, the receivers scatter into theirs, all inside the kernel
            _lib.check(self.lib.dlbb_broadcast_v(self.h, t.data_ptr(), t.numel(), DTYPES[t.dtype], root, C.byref(view),
                                                 _stream(self.device)), "broadcast (strided)")
            return t
        _lib.check(self.lib.dlbb_broadcast(self.h, t.data_ptr(), t.numel(), DTYPES[t.dtype], root,
                                           _stream(self.device)), "broadcast")
        return t

    def reduce(self, t: torch.Tensor, root: int = 0, op: str = "sum", out: torch.Tensor | None = None):
        out = t if out is None else out
        assert t.is_contiguous() and out.is_contiguous()
        _lib.check(self.lib.dlbb_reduce(self.h, t.data_ptr(), out.data_ptr(), t.numel(), DTYPES[t.dtype], OPS[op], root,
                                        _stream(self.device)), "reduce")
        return out

    def gather(self, inp: torch.Tensor, out: torch.Tensor | None, root: int = 0):
        assert inp.is_contiguous()
        optr = out.data_ptr() if (out is not None and self.rank == root) else None
        _lib.check(self.lib.dlbb_gather(self.h, inp.data_ptr(), optr, inp.numel(), DTYPES[inp.dtype], root,
                                        _stream(self.device)), "gather")
        return out

    def scatter(self, inp: torch.Tensor | None, out: torch.Tensor, root: int = 0):
        assert out.is_contiguous()
        iptr = inp.data_ptr() if (inp is not None and self.rank == root) else None
        _lib.check(self.lib.dlbb_scatter(self.h, iptr, out.data_ptr(), out.numel(), DTYPES[out.dtype], root,
                                         _stream(self.device)), "scatter")
        return out

    def all_to_all(self, inp: torch.Tensor, out: torch.Tensor):
        assert inp.is_contiguous() and out.is_contiguous() and inp.numel() == out.numel()
        assert inp.numel() % self.world_size == 0
        _lib.check(self.lib.dlbb_alltoall(self.h, inp.data_ptr(), out.data_ptr(), inp.numel() // self.world_size,
                                          DTYPES[inp.dtype], _stream(self.device)), "alltoall")
        return out

    def sendrecv(self, send: torch.Tensor | None, send_peer: int, recv: torch.Tensor | None, recv_peer: int):
        t = send if send is not None else recv
        _lib.check(self.lib.dlbb_sendrecv(self.h, send.data_ptr() if send is not None else None,
                                          send.numel() if send is not None else 0, send_peer if send is not None else -1,
                                          recv.data_ptr() if recv is not None else None,
                                          recv.numel() if recv is not None else 0, recv_peer if recv is not None else -1,
                                          DTYPES[t.dtype], _stream(self.device)), "sendrecv")
        return recv

    def send(self, t: torch.Tensor, peer: int):
        return self.sendrecv(t, peer, None, -1)

    def recv(self, t: torch.Tensor, peer: int):
        return self.sendrecv(None, -1, t, peer)

    # ------------------------------------------------------------------ fused GEMM + all-reduce (TP row-parallel)
    def alloc_matrix(self, rows: int, cols: int) -> "SymmMatrix":
        """Collective.  A bf16 [rows, cols] matrix in symmetric memory (peers can read / write it)."""
        t, seg = self.alloc(rows * cols, torch.bfloat16)
        return SymmMatrix(t.view(rows, cols), seg, 0)

    def gemm_allreduce(self, a: torch.Tensor, b: torch.Tensor, part: "SymmMatrix", out: "SymmMatrix",
                       residual: torch.Tensor | None = None, transpose_a: bool = False, transpose_b: bool = False,
                       use_multicast: bool | None = None) -> torch.Tensor:
        """out = all_reduce_sum(op(a) @ op(b)) + residual, one kernel (csrc/collectives/gemm_allreduce.cu)."""
        M, K = (a.shape[1], a.shape[0]) if transpose_a else (a.shape[0], a.shape[1])
        N = b.shape[0] if transpose_b else b.shape[1]
        assert tuple(out.tensor.shape) == (M, N) and part.tensor.numel() >= M * N
        assert a.stride(-1) == 1 and b.stride(-1) == 1 and a.dtype == torch.bfloat16
        if use_multicast is None:
            use_multicast = os.environ.get("DLBB_TP_MULTICAST", "1") == "1"
        lda = a.stride(0) if a.shape[0] > 1 else a.shape[1]
        ldb = b.stride(0) if b.shape[0] > 1 else b.shape[1]
        _lib.check(self.lib.dlbb_gemm_allreduce_bf16(
            self.h, a.data_ptr(), b.data_ptr(), M, N, K, lda, ldb, 1 if transpose_a else 0, 0 if transpose_b else 1,
            part.seg, part.elem_off, out.seg, out.elem_off, residual.data_ptr() if residual is not None else None,
            residual.stride(0) if residual is not None else 0, 1 if use_multicast else 0, _stream(self.device)),
            "gemm_allreduce")
        return out.tensor

    # ------------------------------------------------------------------ data-parallel fused step
    def alloc_dp_buffers(self, numel: int) -> tuple[torch.Tensor, torch.Tensor]:
        """Symmetric bf16 parameter and gradient buffers for the DP engine."""
        params, pseg = self.alloc(numel, torch.bfloat16)
        grads, gseg = self.alloc(numel, torch.bfloat16)
        self._dp = (pseg, gseg, params, grads)
        return params, grads

    def register_dp_buffers(self, params: torch.Tensor, grads: torch.Tensor) -> None:
        if self._dp is None or self._dp[2].data_ptr() != params.data_ptr() or self._dp[3].data_ptr() != grads.data_ptr():
            raise RuntimeError("data-parallel buffers must come from SymmComm.alloc_dp_buffers (symmetric memory)")

    def fused_reduce_adam(self, elem_off: int, n: int, m: torch.Tensor, v: torch.Tensor, master: torch.Tensor | None, *,
                          lr: float, beta1: float, beta2: float, eps: float, weight_decay: float, grad_scale: float,
                          step: int, use_multicast: bool | None = None) -> None:
        pseg, gseg = self._dp[0], self._dp[1]
        if use_multicast is None:
            use_multicast = os.environ.get("DLBB_DP_MULTICAST", "1") == "1"
        _lib.check(self.lib.dlbb_fused_reduce_adam_bf16(
            self.h, gseg, pseg, elem_off, n, m.data_ptr(), v.data_ptr(), master.data_ptr() if master is not None else None,
            lr, beta1, beta2, eps, weight_decay, grad_scale, step, 1 if use_multicast else 0, _stream(self.device)),
            "fused_reduce_adam")


class SymmMatrix:
    """A tensor living in a symmetric segment plus where it lives (segment id, element offset)."""

    __slots__ = ("tensor", "seg", "elem_off")

    def __init__(self, tensor: torch.Tensor, seg: int, elem_off: int):
        self.tensor, self.seg, self.elem_off = tensor, seg, elem_off
