"""DeepSpeed + Gloo 1-D collective sweep - the copy of collectives/1d/dsgloo.py that the reference also keeps one level up (the entry point BASELINE.json names).

1-D collective sweep through ``deepspeed.comm`` on backend ``"gloo"``: fp16 tensors of 256 / 16 Ki /
256 Ki / 4 Mi elements, {10, 100} warm-up / measured iterations, ``dist.barrier()`` + perf_counter around
each call (inputs of in-place ops are cloned outside the timed region), timings gathered to rank 0 as a
float64 tensor, ``deepspeed_gloo_<op>_ranks<P>_<size>.json`` (ref collectives/1d/dsgloo.py).

Run:  deepspeed --master_port 29501 --include localhost:0,1 dsgloo.py
"""
import os
import sys

_ROOT = os.path.abspath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
for _p in (_ROOT, os.path.join(_ROOT, "compat")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

from distributed_llm_backend_benchmark_b200.bench import sweep  # noqa: E402

BACKEND = "gloo"
IMPLEMENTATION_NAME = "deepspeed_gloo"
RANK_COUNTS = [1, 2, 4, 8, 16, 32, 56]
DATA_SIZES = dict(sweep.DATA_SIZES_1D)
OPERATIONS = ['allreduce', 'allgather', 'broadcast', 'gather', 'scatter', 'reduce', 'alltoall']
WARMUP_ITERATIONS = 10
MEASUREMENT_ITERATIONS = 100
OUTPUT_DIR = "results/dsgloo"


def init_deepspeed():
    import deepspeed
    import deepspeed.comm as dist
    deepspeed.init_distributed(dist_backend=BACKEND)
    return dist.get_rank(), dist.get_world_size()


def _make(op):
    def bench(rank, world_size, data, iterations):
        import deepspeed.comm as dist
        return getattr(sweep.TorchTransport(dist, rank, world_size, data.device), op)(data, iterations)[0]
    bench.__name__ = f"benchmark_{op}"
    return bench


benchmark_allreduce, benchmark_allgather, benchmark_broadcast, benchmark_gather = map(
    _make, ("allreduce", "allgather", "broadcast", "gather"))
benchmark_scatter, benchmark_reduce, benchmark_alltoall, benchmark_sendrecv = map(
    _make, ("scatter", "reduce", "alltoall", "sendrecv"))


def run_benchmark():
    sweep.run_sweep_1d(sweep.SweepSpec(
        flavour="deepspeed", implementation=IMPLEMENTATION_NAME, backend=BACKEND, rank_counts=RANK_COUNTS,
        operations=OPERATIONS, data_sizes=DATA_SIZES, warmup=WARMUP_ITERATIONS, iterations=MEASUREMENT_ITERATIONS,
        output_dir=OUTPUT_DIR))


if __name__ == "__main__":
    try:
        run_benchmark()
    except Exception as e:  # noqa: BLE001
        import traceback
        print(f"\nError: {e}")
        traceback.print_exc()
        sys.exit(1)
