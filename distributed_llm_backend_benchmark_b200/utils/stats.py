"""Post-processing of sweep results (ref collectives/1d/stats.py, collectives/3d/stats.py).

The numbers and file formats of the reference are reproduced exactly so its published
``_stats.json`` / CSV files can serve as golden data (tests/test_stats.py):

* 1-D: per file mean / median / min / max / std / p95 / p99 in microseconds over all
  ``timings[rank][iter]``, load imbalance = (max rank mean - mean of rank means) / mean of rank means
  in percent, per-rank means, and the reference's "bandwidth" = elements x 2 B x ranks / slowest single
  iteration / 2^30 (one formula for every operation, ref 1d/stats.py:77-129,179) -> ``<stem>_stats.json``
  and ``benchmark_statistics.csv``;
* 3-D: mean / median / min / max in milliseconds -> ``<stem>_stats.json``, a standard CSV sorted by
  operation, ranks, hidden, seq, batch and a transposed CSV (metrics as rows, metadata rows below).

On top of that every record gets NCCL-tests style ``algbw_gbps`` / ``busbw_gbps`` (decimal GB/s, from the
median time; device-timed when the result file carries ``timings_device``) - the figures this project
reports.
"""
from __future__ import annotations

import csv
import json
from pathlib import Path

import numpy as np

CSV_1D = ["mpi_implementation", "operation", "num_ranks", "data_size_name", "num_elements", "mean_time_us",
          "median_time_us", "min_time_us", "max_time_us", "std_dev_us", "p95_time_us", "p99_time_us",
          "load_imbalance_percent", "bandwidth_gbps"]
CSV_3D = ["implementation", "operation", "num_ranks", "hidden_dim", "seq_len", "batch", "tensor_size_mb",
          "num_elements", "mean_time_ms", "median_time_ms", "min_time_ms", "max_time_ms"]
KNOWN_OPS = ("allreduce", "allgather", "broadcast", "gather", "scatter", "reduce", "alltoall", "sendrecv",
             "reducescatter")


def busbw_factor(operation: str, ranks: int) -> float:
    """busbw = algbw x factor (NCCL-tests convention)."""
    if ranks <= 1:
        return 1.0
    if operation == "allreduce":
        return 2.0 * (ranks - 1) / ranks
    if operation in ("allgather", "reducescatter", "alltoall", "gather", "scatter"):
        return (ranks - 1) / ranks
    return 1.0


def bus_bandwidth(operation: str, ranks: int, message_bytes: float, seconds: float) -> tuple[float, float]:
    """(algbw, busbw) in decimal GB/s for a per-rank message of `message_bytes`."""
    if seconds <= 0:
        return 0.0, 0.0
    alg = message_bytes / seconds / 1e9
    if operation in ("allgather", "gather"):      # message_bytes is the per-rank contribution
        alg *= ranks
    return alg, alg * busbw_factor(operation, ranks)


def calculate_statistics_1d(timings_2d) -> dict:
    t = np.array(timings_2d)
    rank_means = np.mean(t, axis=1)
    flat = t.flatten()
    mean_of_means = np.mean(rank_means)
    imbalance = ((np.max(rank_means) - mean_of_means) / mean_of_means) * 100 if mean_of_means > 0 else 0.0
    return {
        "mean_time_us": np.mean(flat) * 1e6,
        "median_time_us": np.median(flat) * 1e6,
        "min_time_us": np.min(flat) * 1e6,
        "max_time_us": np.max(flat) * 1e6,
        "std_dev_us": np.std(flat) * 1e6,
        "p95_time_us": np.percentile(flat, 95) * 1e6,
        "p99_time_us": np.percentile(flat, 99) * 1e6,
        "load_imbalance_percent": imbalance,
        "per_rank_means_us": (rank_means * 1e6).tolist(),
    }


def reference_bandwidth_1d(num_elements: int, time_seconds: float, operation: str, num_ranks: int):
    """The reference's aggregate GiB/s figure (fp16 element size hard-coded, same for every op)."""
    if operation not in KNOWN_OPS or time_seconds <= 0:
        return None
    return (num_elements * 2 * num_ranks / time_seconds) / (1024 ** 3)


def _extra_bw(data: dict, operation: str, ranks: int, message_bytes: float) -> dict:
    dev = data.get("timings_device")
    src = dev if dev else data["timings"]
    med = float(np.median(np.max(np.array(src), axis=0)))   # per iteration: slowest rank, then median
    alg, bus = bus_bandwidth(operation, ranks, message_bytes, med)
    return {"timing_source": "device" if dev else "host", "median_max_over_ranks_us": med * 1e6,
            "algbw_gbps": alg, "busbw_gbps": bus}


def process_1d(input_dir, output_dir, output_csv: str = "benchmark_statistics.csv", verbose: bool = True) -> list[dict]:
    out = Path(output_dir)
    out.mkdir(parents=True, exist_ok=True)
    files = sorted(Path(input_dir).glob("*.json"))
    if not files:
        print(f"No JSON files found in '{input_dir}' directory")
        return []
    results = []
    for f in files:
        if verbose:
            print(f"Processing: {f.name}")
        try:
            data = json.loads(f.read_text())
            stats = calculate_statistics_1d(data["timings"])
            stats["bandwidth_gbps"] = reference_bandwidth_1d(data["num_elements"], stats["max_time_us"] / 1e6,
                                                             data["operation"], data["num_ranks"])
            rec = {"mpi_implementation": data.get("mpi_implementation") or data.get("implementation") or "unknown",
                   "operation": data["operation"], "num_ranks": data["num_ranks"],
                   "data_size_name": data["data_size_name"], "num_elements": data["num_elements"],
                   "dtype": data["dtype"], **stats}
            extra = _extra_bw(data, data["operation"], data["num_ranks"],
                              data["num_elements"] * data.get("element_size", 2))
            (out / (f.stem + "_stats.json")).write_text(json.dumps({**rec, **extra}, indent=2))
            results.append({**rec, **extra})
        except Exception as e:  # one bad file must not stop the batch
            print(f"  ERROR processing {f.name}: {e}")
    if results:
        with open(out / output_csv, "w", newline="") as fh:
            w = csv.DictWriter(fh, fieldnames=CSV_1D, extrasaction="ignore")
            w.writeheader()
            for r in results:
                w.writerow(r)
        with open(out / output_csv.replace(".csv", "_busbw.csv"), "w", newline="") as fh:
            cols = ["mpi_implementation", "operation", "num_ranks", "data_size_name", "num_elements", "timing_source",
                    "median_max_over_ranks_us", "algbw_gbps", "busbw_gbps"]
            w = csv.DictWriter(fh, fieldnames=cols, extrasaction="ignore")
            w.writeheader()
            for r in results:
                w.writerow(r)
        if verbose:
            print(f"Consolidated CSV saved: {out / output_csv}")
            print(f"Total configurations processed: {len(results)}")
    return results


def calculate_statistics_3d(timings_2d) -> dict:
    flat = np.array(timings_2d).flatten()
    return {"mean_time_ms": np.mean(flat) * 1e3, "median_time_ms": np.median(flat) * 1e3,
            "min_time_ms": np.min(flat) * 1e3, "max_time_ms": np.max(flat) * 1e3}


def _sort_key_3d(r):
    return (r["operation"], r["num_ranks"], r["hidden_dim"], r["seq_len"], r["batch"])


def process_3d(input_dir, output_dir, implementation: str, verbose: bool = True) -> list[dict]:
    out = Path(output_dir)
    out.mkdir(parents=True, exist_ok=True)
    files = sorted(Path(input_dir).glob("*.json"))
    if not files:
        print(f"No JSON files found in {input_dir}")
        return []
    results = []
    for f in files:
        if verbose:
            print(f"Processing: {f.name}")
        try:
            data = json.loads(f.read_text())
            shape = data["tensor_shape"]
            rec = {"implementation": data.get("mpi_implementation") or data.get("implementation"),
                   "operation": data["operation"], "num_ranks": data["num_ranks"], "hidden_dim": shape["hidden_dim"],
                   "seq_len": shape["seq_len"], "batch": shape["batch"],
                   "tensor_size_mb": round(data["tensor_size_mb"], 4), "num_elements": data["num_elements"],
                   **calculate_statistics_3d(data["timings"])}
            extra = _extra_bw(data, data["operation"], data["num_ranks"], data["tensor_size_bytes"])
            (out / (f.stem + "_stats.json")).write_text(json.dumps({**rec, **extra}, indent=2))
            results.append({**rec, **extra})
        except Exception as e:
            print(f"  ERROR processing {f.name}: {e}")
    if not results:
        return results
    ordered = sorted(results, key=_sort_key_3d)
    with open(out / f"benchmark_statistics_3d_{implementation}_standard.csv", "w", newline="") as fh:
        w = csv.DictWriter(fh, fieldnames=CSV_3D, extrasaction="ignore")
        w.writeheader()
        for r in ordered:
            w.writerow(r)
    with open(out / f"benchmark_statistics_3d_{implementation}_busbw.csv", "w", newline="") as fh:
        cols = ["implementation", "operation", "num_ranks", "hidden_dim", "seq_len", "batch", "tensor_size_mb",
                "timing_source", "median_max_over_ranks_us", "algbw_gbps", "busbw_gbps"]
        w = csv.DictWriter(fh, fieldnames=cols, extrasaction="ignore")
        w.writeheader()
        for r in ordered:
            w.writerow(r)
    with open(out / f"benchmark_statistics_3d_{implementation}_transpose.csv", "w", newline="") as fh:
        w = csv.writer(fh)
        ids = [f"{r['operation']}_r{r['num_ranks']}_h{r['hidden_dim']}_s{r['seq_len']}_b{r['batch']}" for r in ordered]
        w.writerow(["Metric"] + ids)
        for metric in ("mean_time_ms", "median_time_ms", "min_time_ms", "max_time_ms"):
            w.writerow([metric] + [r[metric] for r in ordered])
        w.writerow([])
        w.writerow(["--- Metadata ---"])
        for key in ("operation", "num_ranks", "hidden_dim", "seq_len", "batch", "tensor_size_mb"):
            w.writerow([key] + [r[key] for r in ordered])
    if verbose:
        print(f"Processed {len(results)} files successfully -> {out}")
    return results
