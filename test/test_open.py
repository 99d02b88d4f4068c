#!/usr/bin/env python3
"""MPI smoke test through the mpi4py facade (covers what the reference's test/test_open.py covers):
point-to-point of Python objects, the object collectives, the numpy buffer collectives, a staggered
barrier, a non-blocking ring and the MAX / MIN / PROD reductions.

    mpirun -n 2 python test/test_open.py
Any failure aborts the whole job with a non-zero exit code."""
import os
import sys
import time

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (_ROOT, os.path.join(_ROOT, "compat")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402
from mpi4py import MPI  # noqa: E402

comm = MPI.COMM_WORLD
rank, size = comm.Get_rank(), comm.Get_size()
CHECKS = []


def check(fn):
    CHECKS.append(fn)
    return fn


def say(msg):
    print(f"[rank {rank}] {msg}", flush=True)


@check
def environment():
    if rank == 0:
        say(f"{MPI.Get_library_version()} | MPI standard {MPI.Get_version()} | host {MPI.Get_processor_name()}")
    assert 0 <= rank < size


@check
def point_to_point_objects():
    if size < 2:
        return
    if rank == 0:
        comm.send({"msg": "ping", "payload": [1, 2, 3], "from": 0}, dest=1, tag=11)
        back = comm.recv(source=1, tag=22)
        assert back == {"msg": "pong", "sum": 6}
    elif rank == 1:
        got = comm.recv(source=0, tag=11)
        assert got["msg"] == "ping" and got["from"] == 0
        comm.send({"msg": "pong", "sum": sum(got["payload"])}, dest=0, tag=22)


@check
def broadcast_object():
    cfg = {"lr": 0.001, "layers": [64, 128], "name": "cfg"} if rank == 0 else None
    cfg = comm.bcast(cfg, root=0)
    assert cfg["layers"] == [64, 128] and cfg["name"] == "cfg"


@check
def scatter_gather_allgather():
    part = comm.scatter([{"chunk": i, "sq": i * i} for i in range(size)] if rank == 0 else None, root=0)
    assert part == {"chunk": rank, "sq": rank * rank}
    gathered = comm.gather(rank * 10, root=0)
    assert gathered == ([r * 10 for r in range(size)] if rank == 0 else None)
    assert comm.allgather(f"r{rank}") == [f"r{r}" for r in range(size)]


@check
def reductions():
    total = comm.reduce(rank + 1, op=MPI.SUM, root=0)
    assert total == (size * (size + 1) // 2 if rank == 0 else None)
    assert comm.allreduce(rank + 1, op=MPI.SUM) == size * (size + 1) // 2
    assert comm.allreduce(rank, op=MPI.MAX) == size - 1
    assert comm.allreduce(rank, op=MPI.MIN) == 0
    want = 1
    for r in range(size):
        want *= r + 1
    assert comm.allreduce(rank + 1, op=MPI.PROD) == want


@check
def numpy_buffers():
    buf = np.arange(100, dtype=np.float64) if rank == 0 else np.empty(100, dtype=np.float64)
    comm.Bcast(buf, root=0)
    assert np.array_equal(buf, np.arange(100, dtype=np.float64))
    send = np.full(1000, rank + 1, dtype=np.float32)
    recv = np.empty_like(send)
    comm.Allreduce(send, recv, op=MPI.SUM)
    assert np.allclose(recv, size * (size + 1) / 2)
    half = np.full(64, rank + 1, dtype=np.float16)
    out = np.empty_like(half)
    comm.Allreduce(half, out, op=MPI.MAX)
    assert np.all(out == size)


@check
def staggered_barrier():
    time.sleep(0.05 * rank)
    t0 = MPI.Wtime()
    comm.Barrier()
    assert MPI.Wtime() - t0 >= 0.0


@check
def nonblocking_ring():
    if size < 2:
        return
    nxt, prv = (rank + 1) % size, (rank - 1) % size
    sreq = comm.isend({"token": rank}, dest=nxt, tag=5)
    rreq = comm.irecv(source=prv, tag=5)
    got = rreq.wait()
    sreq.wait()
    assert got == {"token": prv}
    a, b = np.full(256, rank, dtype=np.float16), np.empty(256, dtype=np.float16)
    s = comm.Isend(a, dest=nxt)
    r = comm.Irecv(b, source=prv)
    s.Wait()
    r.Wait()
    assert np.all(b == prv)


def main():
    try:
        for fn in CHECKS:
            fn()
            comm.Barrier()
            if rank == 0:
                say(f"ok  {fn.__name__}")
    except Exception as e:  # noqa: BLE001
        import traceback
        traceback.print_exc()
        say(f"FAILED: {e}")
        comm.Abort(1)
    if rank == 0:
        say(f"all {len(CHECKS)} MPI checks passed on {size} rank(s)")


if __name__ == "__main__":
    main()
