"""mpi4py-compatible facade over the dlbb communication backend (see MPI.py).

Only what the reference uses is provided: ``from mpi4py import MPI`` with
COMM_WORLD, the buffer API (Allreduce, Bcast, Isend/Irecv, ...), the pickle-based object API
(allreduce, allgather, bcast, gather, scatter, reduce, alltoall, send/recv, isend/irecv) and the
reduction operators SUM / MAX / MIN / PROD."""
__version__ = "4.0.0+dlbb"


def get_config():
    return {"backend": "dlbb", "launcher": "b200run"}
