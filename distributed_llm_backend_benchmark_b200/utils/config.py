"""Experiment configuration (ref utils.py:90-102,154-169 and config/baseline_config.yaml).

The YAML schema is the reference's: ``experiment{name,output_dir}``, ``model{size,hidden_size,
num_layers,num_heads,ffn_intermediate}``, ``parallelism{world_size,cores_per_rank}``,
``input{batch_size,sequence_length,seed}``, ``execution{warmup_iterations,benchmark_iterations}``,
``system{omp_num_threads,mkl_num_threads}``.  On top of it, ``DLBB_*`` environment variables
override single keys so the sweeps can run at 1/2/4/8 GPUs without editing files.
"""
from __future__ import annotations

import copy
import os
from pathlib import Path

import yaml

DEFAULT_CONFIG = {
    "experiment": {"name": "baseline_7b_world4", "output_dir": "results"},
    "model": {"size": "7B", "hidden_size": 4096, "num_layers": 32, "num_heads": 32, "ffn_intermediate": 16384},
    "parallelism": {"world_size": 4, "cores_per_rank": 14},
    "input": {"batch_size": 8, "sequence_length": 512, "seed": 42},
    "execution": {"warmup_iterations": 5, "benchmark_iterations": 10},
    "system": {"omp_num_threads": 14, "mkl_num_threads": 14},
}

# env var -> (section, key, type)
ENV_OVERRIDES = {
    "DLBB_WORLD_SIZE": ("parallelism", "world_size", int),
    "DLBB_BATCH_SIZE": ("input", "batch_size", int),
    "DLBB_SEQ_LEN": ("input", "sequence_length", int),
    "DLBB_SEED": ("input", "seed", int),
    "DLBB_WARMUP": ("execution", "warmup_iterations", int),
    "DLBB_ITERS": ("execution", "benchmark_iterations", int),
    "DLBB_NUM_LAYERS": ("model", "num_layers", int),
    "DLBB_OUTPUT_DIR": ("experiment", "output_dir", str),
}


def load_config(config_path: str | os.PathLike | None) -> dict:
    """Parse the YAML file; missing sections fall back to the reference's baseline values."""
    cfg = copy.deepcopy(DEFAULT_CONFIG)
    if config_path is not None:
        with open(Path(config_path), "r") as fh:
            loaded = yaml.safe_load(fh) or {}
        for section, body in loaded.items():
            if isinstance(body, dict) and isinstance(cfg.get(section), dict):
                cfg[section].update(body)
            else:
                cfg[section] = body
    for env, (section, key, typ) in ENV_OVERRIDES.items():
        if env in os.environ:
            cfg.setdefault(section, {})[key] = typ(os.environ[env])
    return cfg


def setup_environment(config: dict) -> None:
    """Export the host threading knobs of the ``system`` section (ref utils.py:154-169)."""
    system = config.get("system", {})
    for key, env in (("omp_num_threads", "OMP_NUM_THREADS"), ("mkl_num_threads", "MKL_NUM_THREADS")):
        if key in system:
            os.environ[env] = str(system[key])


def validate_tp(config: dict, world_size: int) -> None:
    """The reference shards with bare integer division (models.py:33,66); make the silent
    truncation an error instead."""
    m = config["model"]
    for name, dim in (("3*hidden_size", 3 * m["hidden_size"]), ("hidden_size", m["hidden_size"]),
                      ("ffn_intermediate", m["ffn_intermediate"])):
        if dim % world_size:
            raise ValueError(f"{name}={dim} is not divisible by the tensor-parallel size {world_size}")
