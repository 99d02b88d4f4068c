"""``mpi4py.MPI`` on top of torch.distributed + the dlbb backend.

The reference talks to MPI in two ways:
  * buffer API, upper-case methods on numpy arrays - ``comm.Allreduce(a, b, op=MPI.SUM)``
    (models.py:95), ``comm.Bcast`` (collectives/1d/openmpi.py:98), ``Isend/Irecv`` rings (:189-193);
  * object API, lower-case methods that pickle arbitrary Python objects - ``comm.allreduce(arr)``
    (collectives/1d/openmpi.py:63), ``gather`` / ``scatter`` / ``alltoall`` of lists, dict ping-pong
    (test/test_open.py:35-63), reductions with MAX / MIN / PROD (test/test_open.py:248-271).
Both are mapped onto torch.distributed collectives: numpy buffers are wrapped zero-copy as CPU
tensors (the dlbb backend moves them through the rank's GPU and reduces them with its own kernels;
on a CPU-only host the same calls run over Gloo), objects travel as uint8 tensors and are combined
in Python, exactly like mpi4py's pickle path.

Rank and size come from the launcher's environment (bin/mpirun -> b200run sets OMPI_COMM_WORLD_*,
PMI_* and the torch-style RANK / WORLD_SIZE / MASTER_*).  Without a launcher the world has one rank.
"""
from __future__ import annotations

import os
import pickle
import socket
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ANY_SOURCE = -1
ANY_TAG = -1
PROC_NULL = -2
IN_PLACE = "MPI_IN_PLACE"
UNDEFINED = -32766


class Op:
    def __init__(self, name, py, torch_op):
        self.name, self._py, self.torch_op = name, py, torch_op

    def __call__(self, a, b):
        return self._py(a, b)

    def __repr__(self):
        return f"<mpi4py.MPI.Op {self.name}>"


def _max(a, b):
    return np.maximum(a, b) if isinstance(a, np.ndarray) or isinstance(b, np.ndarray) else max(a, b)


def _min(a, b):
    return np.minimum(a, b) if isinstance(a, np.ndarray) or isinstance(b, np.ndarray) else min(a, b)


SUM = Op("SUM", lambda a, b: a + b, dist.ReduceOp.SUM)
PROD = Op("PROD", lambda a, b: a * b, dist.ReduceOp.PRODUCT)
MAX = Op("MAX", _max, dist.ReduceOp.MAX)
MIN = Op("MIN", _min, dist.ReduceOp.MIN)
LAND = Op("LAND", lambda a, b: bool(a) and bool(b), None)
LOR = Op("LOR", lambda a, b: bool(a) or bool(b), None)

# numpy dtypes torch cannot wrap natively are sent as raw bytes
_NATIVE = {np.dtype(t) for t in (np.float16, np.float32, np.float64, np.int32, np.int64, np.uint8, np.int8, np.int16, np.bool_)}


def _env_int(*names, default):
    for n in names:
        if n in os.environ:
            return int(os.environ[n])
    return default


class Status:
    def __init__(self):
        self.source, self.tag, self.count = ANY_SOURCE, ANY_TAG, 0

    def Get_source(self):
        return self.source

    def Get_tag(self):
        return self.tag

    def Get_count(self, datatype=None):
        return self.count


class Request:
    """Completion handle; ``wait()`` returns the received object for irecv (mpi4py semantics)."""

    def __init__(self, finish=None):
        self._finish, self._done, self._value = finish, finish is None, None

    def Wait(self, status=None):
        self.wait(status)
        return True

    def wait(self, status=None):
        if not self._done:
            self._value = self._finish()
            self._done = True
        return self._value

    def Test(self, status=None):
        return self._done

    def test(self, status=None):
        return (self._done, self._value)

    @staticmethod
    def Waitall(requests, statuses=None):
        for r in requests:
            r.wait()
        return True

    waitall = Waitall


def _as_tensor(buf):
    """numpy array (or [array, count, type] spec, or torch tensor) -> flat torch view sharing memory."""
    if isinstance(buf, (list, tuple)):
        buf = buf[0]
    if isinstance(buf, torch.Tensor):
        return buf.view(-1) if buf.is_contiguous() else buf
    arr = np.asarray(buf)
    if not arr.flags["C_CONTIGUOUS"]:
        raise ValueError("mpi4py shim: buffer-API arrays must be C-contiguous")
    if arr.dtype in _NATIVE:
        return torch.from_numpy(arr.reshape(-1))
    return torch.from_numpy(arr.reshape(-1).view(np.uint8))


class Comm:
    def __init__(self):
        self._rank = _env_int("RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", default=0)
        self._size = _env_int("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", default=1)
        self._ready = False

    # ------------------------------------------------------------------ plumbing
    def _init(self):
        if self._ready or self._size == 1:
            self._ready = True
            return
        if not dist.is_initialized():
            from distributed_llm_backend_benchmark_b200.parallel import backend as be

            os.environ.setdefault("RANK", str(self._rank))
            os.environ.setdefault("WORLD_SIZE", str(self._size))
            be.init_process_group(os.environ.get("DLBB_MPI_BACKEND"), rank=self._rank, world_size=self._size)
        self._ready = True

    def Get_rank(self):
        return self._rank

    def Get_size(self):
        return self._size

    rank = property(Get_rank)
    size = property(Get_size)

    def Get_name(self):
        return "MPI_COMM_WORLD"

    def Dup(self):
        return self

    def Free(self):
        pass

    def Abort(self, errorcode=1):
        sys.stdout.flush()
        sys.stderr.write(f"MPI_ABORT was invoked on rank {self._rank} with errorcode {errorcode}\n")
        sys.stderr.flush()
        os._exit(int(errorcode) or 1)  # the launcher tears the rest of the job down

    def Barrier(self):
        self._init()
        if self._size > 1:
            dist.barrier()

    barrier = Barrier

    # ------------------------------------------------------------------ buffer API
    def Allreduce(self, sendbuf, recvbuf, op=SUM):
        self._init()
        out = _as_tensor(recvbuf)
        if sendbuf is not IN_PLACE:
            out.copy_(_as_tensor(sendbuf))
        if self._size > 1:
            dist.all_reduce(out, op=op.torch_op)

    def Reduce(self, sendbuf, recvbuf, op=SUM, root=0):
        self._init()
        src = _as_tensor(recvbuf if sendbuf is IN_PLACE else sendbuf)
        tmp = src.clone()
        if self._size > 1:
            dist.reduce(tmp, dst=root, op=op.torch_op)
        if self._rank == root:
            _as_tensor(recvbuf).copy_(tmp)

    def Bcast(self, buf, root=0):
        self._init()
        if self._size > 1:
            dist.broadcast(_as_tensor(buf), src=root)

    def Allgather(self, sendbuf, recvbuf):
        self._init()
        out, src = _as_tensor(recvbuf), _as_tensor(sendbuf)
        if self._size == 1:
            out.copy_(src)
        else:
            dist.all_gather_into_tensor(out, src.contiguous())

    def Gather(self, sendbuf, recvbuf, root=0):
        self._init()
        src = _as_tensor(sendbuf)
        if self._size == 1:
            _as_tensor(recvbuf).copy_(src)
            return
        if self._rank == root:
            out = _as_tensor(recvbuf)
            dist.gather(src, list(out.view(self._size, -1).unbind(0)), dst=root)
        else:
            dist.gather(src, None, dst=root)

    def Scatter(self, sendbuf, recvbuf, root=0):
        self._init()
        out = _as_tensor(recvbuf)
        if self._size == 1:
            out.copy_(_as_tensor(sendbuf))
            return
        if self._rank == root:
            dist.scatter(out, list(_as_tensor(sendbuf).view(self._size, -1).unbind(0)), src=root)
        else:
            dist.scatter(out, None, src=root)

    def Alltoall(self, sendbuf, recvbuf):
        self._init()
        out, src = _as_tensor(recvbuf), _as_tensor(sendbuf)
        if self._size == 1:
            out.copy_(src)
        else:
            dist.all_to_all_single(out, src.contiguous())

    def Send(self, buf, dest, tag=0):
        self._init()
        dist.send(_as_tensor(buf), dst=dest)

    def Recv(self, buf, source=ANY_SOURCE, tag=ANY_TAG, status=None):
        self._init()
        dist.recv(_as_tensor(buf), src=source)

    def Isend(self, buf, dest, tag=0):
        self._init()
        w = dist.isend(_as_tensor(buf), dst=dest)
        return Request(lambda: w.wait())

    def Irecv(self, buf, source=ANY_SOURCE, tag=ANY_TAG):
        self._init()
        w = dist.irecv(_as_tensor(buf), src=source)
        return Request(lambda: w.wait())

    def Sendrecv(self, sendbuf, dest, sendtag=0, recvbuf=None, source=ANY_SOURCE, recvtag=ANY_TAG, status=None):
        r = self.Irecv(recvbuf, source)
        s = self.Isend(sendbuf, dest)
        s.wait()
        r.wait()

    # ------------------------------------------------------------------ object API (pickle path)
    @staticmethod
    def _dumps(obj) -> torch.Tensor:
        raw = pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
        return torch.frombuffer(bytearray(raw), dtype=torch.uint8)

    @staticmethod
    def _loads(t: torch.Tensor):
        return pickle.loads(t.numpy().tobytes())

    def _allgather_objs(self, obj):
        """Every rank's object, in rank order (two collectives: sizes, then padded payloads)."""
        if self._size == 1:
            return [obj]
        payload = self._dumps(obj)
        sizes = torch.zeros(self._size, dtype=torch.int64)
        dist.all_gather_into_tensor(sizes, torch.tensor([payload.numel()], dtype=torch.int64))
        width = int(sizes.max())
        padded = torch.zeros(width, dtype=torch.uint8)
        padded[:payload.numel()] = payload
        allb = torch.empty(self._size * width, dtype=torch.uint8)
        dist.all_gather_into_tensor(allb, padded)
        return [self._loads(allb[p * width:p * width + int(sizes[p])]) for p in range(self._size)]

    def bcast(self, obj=None, root=0):
        self._init()
        if self._size == 1:
            return obj
        n = torch.zeros(1, dtype=torch.int64)
        payload = None
        if self._rank == root:
            payload = self._dumps(obj)
            n[0] = payload.numel()
        dist.broadcast(n, src=root)
        if self._rank != root:
            payload = torch.empty(int(n), dtype=torch.uint8)
        dist.broadcast(payload, src=root)
        return obj if self._rank == root else self._loads(payload)

    def allgather(self, obj):
        self._init()
        return self._allgather_objs(obj)

    def gather(self, obj, root=0):
        self._init()
        objs = self._allgather_objs(obj)
        return objs if self._rank == root else None

    def scatter(self, objs=None, root=0):
        self._init()
        if self._size == 1:
            return objs[0]
        if self._rank == root and (objs is None or len(objs) != self._size):
            raise ValueError("scatter: the root must pass one object per rank")
        table = self.bcast(objs if self._rank == root else None, root=root)
        return table[self._rank]

    def alltoall(self, objs):
        self._init()
        if len(objs) != self._size:
            raise ValueError("alltoall: one object per rank expected")
        table = self._allgather_objs(list(objs))
        return [table[p][self._rank] for p in range(self._size)]

    def allreduce(self, obj, op=SUM):
        self._init()
        objs = self._allgather_objs(obj)
        acc = objs[0]
        for o in objs[1:]:
            acc = op(acc, o)
        return acc

    def reduce(self, obj, op=SUM, root=0):
        res = self.allreduce(obj, op)
        return res if self._rank == root else None

    def send(self, obj, dest, tag=0):
        self._init()
        payload = self._dumps(obj)
        dist.send(torch.tensor([payload.numel()], dtype=torch.int64), dst=dest)
        dist.send(payload, dst=dest)

    def recv(self, buf=None, source=ANY_SOURCE, tag=ANY_TAG, status=None):
        self._init()
        if source in (ANY_SOURCE, None):
            raise NotImplementedError("mpi4py shim: recv needs an explicit source rank")
        n = torch.zeros(1, dtype=torch.int64)
        dist.recv(n, src=source)
        payload = torch.empty(int(n), dtype=torch.uint8)
        dist.recv(payload, src=source)
        if status is not None:
            status.source, status.count = source, int(n)
        return self._loads(payload)

    def isend(self, obj, dest, tag=0):
        self._init()
        payload = self._dumps(obj)
        header = torch.tensor([payload.numel()], dtype=torch.int64)
        w1 = dist.isend(header, dst=dest)
        w2 = dist.isend(payload, dst=dest)
        keep = (header, payload)

        def finish():
            w1.wait()
            w2.wait()
            return keep and None

        return Request(finish)

    def irecv(self, buf=None, source=ANY_SOURCE, tag=ANY_TAG):
        self._init()
        return Request(lambda: self.recv(source=source))

    def sendrecv(self, sendobj, dest, sendtag=0, recvbuf=None, source=ANY_SOURCE, recvtag=ANY_TAG, status=None):
        s = self.isend(sendobj, dest)
        r = self.recv(source=source)
        s.wait()
        return r


Intracomm = Comm
COMM_WORLD = Comm()
COMM_SELF = COMM_WORLD if COMM_WORLD.Get_size() == 1 else None

_t0 = time.perf_counter()


def Wtime() -> float:
    return time.perf_counter() - _t0 + 1.0e9 * 0  # seconds since import, monotonic


def Wtick() -> float:
    return 1e-9


def Get_processor_name() -> str:
    return socket.gethostname()


def Get_version():
    return (4, 0)


def Get_library_version() -> str:
    import distributed_llm_backend_benchmark_b200 as pkg

    return (f"dlbb {pkg.__version__}: B200-native MPI facade (NVLink P2P / NVLS collectives, "
            f"torch {torch.__version__})")


def Is_initialized() -> bool:
    return True


def Is_finalized() -> bool:
    return False


def Init():
    pass


def Finalize():
    if dist.is_initialized():
        try:
            dist.barrier()
            dist.destroy_process_group()
        except Exception:
            pass


import atexit as _atexit  # noqa: E402

_atexit.register(Finalize)
