"""Statistics for the 1-D sweep: one ``<stem>_stats.json`` per result file plus
``benchmark_statistics.csv`` (ref collectives/1d/stats.py).  Usage:

    python stats.py                      # INPUT_DIR / OUTPUT_DIR below
    python stats.py results/openmpi stats/openmpi
"""
import os
import sys

_ROOT = os.path.abspath(os.path.join(os.path.dirname(os.path.abspath(__file__)), "../.."))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from distributed_llm_backend_benchmark_b200.utils import stats  # noqa: E402

INPUT_DIR = os.environ.get("DLBB_STATS_INPUT", "results/dsccl")
OUTPUT_DIR = os.environ.get("DLBB_STATS_OUTPUT", "stats/dsccl")
OUTPUT_CSV = "benchmark_statistics.csv"

calculate_statistics = stats.calculate_statistics_1d


def calculate_bandwidth(num_elements, dtype, time_seconds, operation, num_ranks):
    return stats.reference_bandwidth_1d(num_elements, time_seconds, operation, num_ranks)


def process_all_json_files():
    return stats.process_1d(INPUT_DIR, OUTPUT_DIR, OUTPUT_CSV)


if __name__ == "__main__":
    if len(sys.argv) >= 3:
        INPUT_DIR, OUTPUT_DIR = sys.argv[1], sys.argv[2]
    process_all_json_files()
