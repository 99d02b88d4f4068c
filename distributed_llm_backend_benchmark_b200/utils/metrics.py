"""Timing and metrics (ref utils.py:17-87,172-265 and run_mpi.py:193-225).

``MetricsCollector`` keeps the reference's fields (init / warm-up / forward times and the
mean, std, min, max, median, p95, p99 summary).  Host wall-clock ``Timer`` is kept for schema
compatibility; ``DeviceTimer`` adds CUDA-event timing, which is what every reported multi-GPU
number uses (max over ranks).
"""
from __future__ import annotations

import time

import numpy as np


class MetricsCollector:
    def __init__(self, rank: int, world_size: int):
        self.rank = rank
        self.world_size = world_size
        self.metrics = {"init_time": None, "warmup_times": [], "forward_times": [], "forward_times_device": []}

    def record_init_time(self, elapsed: float) -> None:
        self.metrics["init_time"] = elapsed

    def record_warmup_time(self, elapsed: float) -> None:
        self.metrics["warmup_times"].append(elapsed)

    def record_forward_time(self, elapsed: float, device_elapsed: float | None = None) -> None:
        self.metrics["forward_times"].append(elapsed)
        if device_elapsed is not None:
            self.metrics["forward_times_device"].append(device_elapsed)

    @staticmethod
    def _stats(prefix: str, values) -> dict:
        t = np.asarray(values, dtype=np.float64)
        if t.size == 0:
            return {}
        return {
            f"{prefix}_mean": float(t.mean()), f"{prefix}_std": float(t.std()), f"{prefix}_min": float(t.min()),
            f"{prefix}_max": float(t.max()), f"{prefix}_median": float(np.median(t)),
            f"{prefix}_p95": float(np.percentile(t, 95)), f"{prefix}_p99": float(np.percentile(t, 99)),
        }

    def get_summary(self) -> dict:
        out = {"rank": self.rank, "world_size": self.world_size, "init_time": self.metrics["init_time"],
               "num_iterations": len(self.metrics["forward_times"])}
        out.update(self._stats("forward", self.metrics["forward_times"]))
        out.update(self._stats("forward_device", self.metrics["forward_times_device"]))
        return out

    def get_raw_metrics(self) -> dict:
        return self.metrics


class Timer:
    """``with Timer() as t: ...; t.elapsed`` - host seconds."""

    def __init__(self):
        self.start_time = None
        self.elapsed = None

    def __enter__(self):
        self.start_time = time.time()
        return self

    def __exit__(self, *exc):
        self.elapsed = time.time() - self.start_time


class DeviceTimer:
    """CUDA-event timer on the current stream; ``elapsed`` (seconds) is valid after ``sync()``.
    Falls back to the host clock for CPU runs."""

    def __init__(self, device=None):
        import torch

        self._cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        if self._cuda:
            self._e0 = torch.cuda.Event(enable_timing=True)
            self._e1 = torch.cuda.Event(enable_timing=True)
        self.elapsed = None

    def __enter__(self):
        if self._cuda:
            self._e0.record()
        else:
            self._t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        if self._cuda:
            self._e1.record()
        else:
            self.elapsed = time.perf_counter() - self._t0

    def sync(self) -> float:
        if self._cuda:
            self._e1.synchronize()
            self.elapsed = self._e0.elapsed_time(self._e1) / 1e3
        return self.elapsed


def rank_statistics(per_rank_means) -> dict:
    """Cross-rank spread exactly as run_mpi.py:201-212 reports it."""
    arr = np.asarray(per_rank_means, dtype=np.float64)
    mean = float(arr.mean())
    return {
        "forward_mean_per_rank": arr.tolist(),
        "variance_across_ranks": float(arr.var()),
        "coefficient_of_variation": float(arr.std() / mean) if mean else 0.0,
    }


def gather_metrics_from_all_ranks(local_summary: dict, rank: int, world_size: int):
    """torch.distributed flavour of the rank statistics (ref utils.py:172-209)."""
    import torch
    import torch.distributed as dist

    mine = torch.tensor([local_summary["forward_mean"]], dtype=torch.float32)
    if world_size == 1 or not dist.is_initialized():
        return rank_statistics([float(mine)]) if rank == 0 else None
    if rank == 0:
        bucket = [torch.zeros(1) for _ in range(world_size)]
        dist.gather(mine, bucket, dst=0)
        return rank_statistics([float(b) for b in bucket])
    dist.gather(mine, dst=0)
    return None


def print_summary(summary: dict, backend_name: str, rank: int) -> None:
    if rank != 0:
        return
    bar = "=" * 60
    print(f"\n{bar}\nRESULTS - {backend_name}\n{bar}")
    print(f"Initialization time: {summary['init_time']:.4f}s")
    for key, label in (("forward_mean", "mean"), ("forward_std", "std"), ("forward_p95", "p95"),
                       ("forward_p99", "p99")):
        print(f"Forward pass {label + ':':<6} {summary[key]:.4f}s")
    if "forward_device_mean" in summary:
        print(f"Forward (device) mean: {summary['forward_device_mean']:.4f}s")
    print(bar)
