"""Result files and system description (ref utils.py:105-151,268-279)."""
from __future__ import annotations

import json
import platform
from pathlib import Path


def save_json(data, filepath) -> None:
    path = Path(filepath)
    path.parent.mkdir(parents=True, exist_ok=True)
    with open(path, "w") as fh:
        json.dump(data, fh, indent=2)


def load_json(filepath):
    with open(filepath, "r") as fh:
        return json.load(fh)


def save_results(results, output_path, rank: int) -> None:
    """Rank 0 writes; everyone else returns immediately."""
    if rank == 0:
        save_json(results, output_path)
        print(f"\nResults saved to: {output_path}")


def collect_system_info() -> dict:
    """Host description with the reference's keys plus what matters on a GPU box."""
    import psutil
    import torch

    info = {
        "platform": platform.platform(),
        "python_version": platform.python_version(),
        "torch_version": torch.__version__,
        "cpu_count": psutil.cpu_count(logical=False),
        "cpu_count_logical": psutil.cpu_count(logical=True),
        "total_memory_gb": psutil.virtual_memory().total / (1024 ** 3),
    }
    if torch.cuda.is_available():
        p = torch.cuda.get_device_properties(0)
        info.update({"gpu_count": torch.cuda.device_count(), "gpu_name": p.name,
                     "gpu_memory_gb": p.total_memory / (1024 ** 3), "gpu_sm_count": p.multi_processor_count,
                     "gpu_compute_capability": f"{p.major}.{p.minor}", "cuda_version": torch.version.cuda})
    return info
