"""DeepSpeed + 'ccl' flavour of the 3-D tensor collective sweep ('ccl' = the dlbb NVLink backend).

3-D tensor sweep through ``deepspeed.comm`` on backend ``"ccl"``: bf16 ``(batch, seq, hidden)`` tensors
(no up-cast), all-reduce / all-gather / broadcast / gather / reduce, batch x seq x hidden = {1,8,16,32} x
{1,2048,4096,8192} x {2048,4096} (up to 2 GiB), ``deepspeed_oneccl_<op>_ranks<P>_b<b>_s<s>_h<h>.json``
(ref collectives/3d/dsccl.py).  DLBB_SWEEP_STRIDED=1 communicates non-contiguous [b,s,h] views, which
the backend packs inside the collective kernel (BASELINE.json config 3).

Run:  deepspeed --num_gpus <P> --master_port 29502 dsccl.py      (see launch_dsccl.sh)
"""
import os
import sys

_ROOT = os.path.abspath(os.path.join(os.path.dirname(os.path.abspath(__file__)), "../.."))
for _p in (_ROOT, os.path.join(_ROOT, "compat")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

from distributed_llm_backend_benchmark_b200.bench import sweep  # noqa: E402

BACKEND = "ccl"
IMPLEMENTATION_NAME = "deepspeed_oneccl"
RANK_COUNTS = [1, 2, 4, 8, 16]
BATCH_SIZES = [1, 8, 16, 32]
SEQ_LENGTHS = [1, 2048, 4096, 8192]
HIDDEN_DIMS = [2048, 4096]
OPERATIONS = list(sweep.OPS_3D)
WARMUP_ITERATIONS = 10
MEASUREMENT_ITERATIONS = 100
OUTPUT_DIR = "results/dsccl"


def init_deepspeed():
    import deepspeed
    import deepspeed.comm as dist
    deepspeed.init_distributed(dist_backend=BACKEND)
    return dist.get_rank(), dist.get_world_size()


def run_benchmark():
    sweep.run_sweep_3d(sweep.SweepSpec(
        flavour="deepspeed", implementation=IMPLEMENTATION_NAME, backend=BACKEND, rank_counts=RANK_COUNTS,
        operations=OPERATIONS, batch_sizes=BATCH_SIZES, seq_lengths=SEQ_LENGTHS, hidden_dims=HIDDEN_DIMS,
        warmup=WARMUP_ITERATIONS, iterations=MEASUREMENT_ITERATIONS, output_dir=OUTPUT_DIR))


if __name__ == "__main__":
    try:
        run_benchmark()
    except Exception as e:  # noqa: BLE001
        import traceback
        print(f"\nError: {e}")
        traceback.print_exc()
        sys.exit(1)
