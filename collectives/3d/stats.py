"""Statistics for the 3-D sweep, one implementation at a time: per-file ``_stats.json``, a standard CSV
(one row per configuration, sorted operation -> ranks -> hidden -> seq -> batch) and a transposed CSV
(metrics as rows) - ref collectives/3d/stats.py.  Usage:

    python stats.py                          # IMPLEMENTATION below
    python stats.py dsccl [input_dir] [output_dir]
"""
import os
import sys

_ROOT = os.path.abspath(os.path.join(os.path.dirname(os.path.abspath(__file__)), "../.."))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from distributed_llm_backend_benchmark_b200.utils import stats  # noqa: E402

IMPLEMENTATION = os.environ.get("DLBB_STATS_IMPL", "dsccl")
INPUT_DIRS = {name: f"results/{name}" for name in ("openmpi", "intelmpi", "dsgloo", "dsccl", "mpitorchccl")}
OUTPUT_DIR = f"stats/{IMPLEMENTATION}"

calculate_statistics = stats.calculate_statistics_3d


def process_implementation(implementation=None, input_dir=None, output_dir=None):
    impl = implementation or IMPLEMENTATION
    if input_dir is None and impl not in INPUT_DIRS:
        print(f"Error: Unknown implementation '{impl}'.  Available: {list(INPUT_DIRS)}")
        sys.exit(1)
    return stats.process_3d(input_dir or INPUT_DIRS[impl], output_dir or f"stats/{impl}", impl)


if __name__ == "__main__":
    a = sys.argv[1:]
    process_implementation(a[0] if a else None, a[1] if len(a) > 1 else None, a[2] if len(a) > 2 else None)
