"""Intel MPI flavour of the 1-D collective sweep (same engine, different labels and output directory).

1-D collective sweep over the mpi4py API: fp16 vectors of 256 / 16 Ki / 256 Ki / 4 Mi elements, eight
operations, {warm-up, measurement} = {10, 100} iterations, Barrier + MPI.Wtime() around each call,
timings gathered to rank 0 and written as ``intelmpi_<op>_ranks<P>_<size>.json`` (ref
collectives/1d/intelmpi.py).  ``mpi4py`` is the dlbb facade, so the object-API calls pickle on the host
and the buffer-API calls (Bcast, Isend/Irecv) run the native NVLink kernels.

Run:  mpirun -np <P> python intelmpi.py        (P in RANK_COUNTS)
"""
import os
import sys

_ROOT = os.path.abspath(os.path.join(os.path.dirname(os.path.abspath(__file__)), "../.."))
for _p in (_ROOT, os.path.join(_ROOT, "compat")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

from distributed_llm_backend_benchmark_b200.bench import sweep  # noqa: E402

import numpy as np  # noqa: E402

MPI_IMPLEMENTATION = "intelmpi"
DTYPE = np.float16
RANK_COUNTS = [1, 2, 4, 8, 16, 32, 56]
DATA_SIZES = dict(sweep.DATA_SIZES_1D)       # labels are nominal: "1KB" = 256 fp16 = 512 B (as in the reference)
OPERATIONS = list(sweep.OPS_1D)
WARMUP_ITERATIONS = 10
MEASUREMENT_ITERATIONS = 100
OUTPUT_DIR = "results/intelmpi"


def _transport():
    from mpi4py import MPI
    return sweep.MpiTransport(MPI, MPI.COMM_WORLD)


def _make(op):
    def bench(comm, data, iterations):
        """Per-iteration wall times (seconds) of `{}` on this rank.""".format(op)
        from mpi4py import MPI
        return getattr(sweep.MpiTransport(MPI, comm), op)(data, iterations)[0]
    bench.__name__ = f"benchmark_{op}"
    return bench


benchmark_allreduce, benchmark_allgather, benchmark_broadcast, benchmark_gather = map(
    _make, ("allreduce", "allgather", "broadcast", "gather"))
benchmark_scatter, benchmark_reduce, benchmark_alltoall, benchmark_sendrecv = map(
    _make, ("scatter", "reduce", "alltoall", "sendrecv"))


def run_benchmark():
    sweep.run_sweep_1d(sweep.SweepSpec(
        flavour="mpi", implementation=MPI_IMPLEMENTATION, rank_counts=RANK_COUNTS, operations=OPERATIONS,
        data_sizes=DATA_SIZES, warmup=WARMUP_ITERATIONS, iterations=MEASUREMENT_ITERATIONS, output_dir=OUTPUT_DIR,
        exit_on_bad_world=False))


if __name__ == "__main__":
    run_benchmark()
