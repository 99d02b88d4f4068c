"""Intel MPI flavour of the 3-D tensor collective sweep.

3-D tensor sweep over mpi4py: bf16 ``(batch, seq, hidden)`` tensors up-cast to fp32 numpy before they are
sent (numpy has no bf16 - ref collectives/3d/intelmpi.py:43; sizes are still reported as bf16 bytes), five
operations, batch x seq x hidden = {1,8,16,32} x {1,2048,4096,8192} x {2048,4096}, results in
``intelmpi_<op>_ranks<P>_b<b>_s<s>_h<h>.json``.

Run:  mpirun -np <P> python intelmpi.py
"""
import os
import sys

_ROOT = os.path.abspath(os.path.join(os.path.dirname(os.path.abspath(__file__)), "../.."))
for _p in (_ROOT, os.path.join(_ROOT, "compat")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

from distributed_llm_backend_benchmark_b200.bench import sweep  # noqa: E402

MPI_IMPLEMENTATION = "intelmpi"
RANK_COUNTS = [1, 2, 4, 8, 16]
BATCH_SIZES = [1, 8, 16, 32]
SEQ_LENGTHS = [1, 2048, 4096, 8192]
HIDDEN_DIMS = [2048, 4096]
OPERATIONS = list(sweep.OPS_3D)
WARMUP_ITERATIONS = 10
MEASUREMENT_ITERATIONS = 100
OUTPUT_DIR = "results/intelmpi"


def run_benchmark():
    sweep.run_sweep_3d(sweep.SweepSpec(
        flavour="mpi", implementation=MPI_IMPLEMENTATION, rank_counts=RANK_COUNTS, operations=OPERATIONS,
        batch_sizes=BATCH_SIZES, seq_lengths=SEQ_LENGTHS, hidden_dims=HIDDEN_DIMS, warmup=WARMUP_ITERATIONS,
        iterations=MEASUREMENT_ITERATIONS, output_dir=OUTPUT_DIR, exit_on_bad_world=False))


if __name__ == "__main__":
    run_benchmark()
