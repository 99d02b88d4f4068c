#!/usr/bin/env python3
"""Tensor-parallel LLM forward benchmark - the reference's ``run_mpi.py`` entry point on B200.

    mpirun -np 4 python run_mpi.py --config config/baseline_config.yaml --backend openmpi

Same command line, same phases (rank/size from ``mpi4py.MPI.COMM_WORLD``, world-size check against
the YAML, warm-up loop, barrier-bracketed timed loop, per-rank means gathered to rank 0, cross-rank
variance / CV, JSON written to ``{output_dir}/{backend}_{experiment.name}.json`` with the
reference's keys - ref run_mpi.py:52-248).  What differs is below the API: ``mpi4py`` is the dlbb
facade, the model lives on the rank's GPU, the batch is uploaded once, and each timed forward is
measured twice - with the reference's host clock between barriers and with CUDA events on the
device (``forward_device_*`` keys, the numbers this project reports).

``--backend`` only names the output file, as in the reference; ``dlbb`` is accepted as well.
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import torch
from mpi4py import MPI

from data_gen import create_dataset_from_config
from models import create_model_from_config
from utils import (DeviceTimer, MetricsCollector, Timer, collect_system_info, load_config, print_summary,
                   rank_statistics, save_results, setup_environment)

RULE = "=" * 60


def initialize_mpi_backend():
    comm = MPI.COMM_WORLD
    return comm.Get_rank(), comm.Get_size(), comm


def cleanup_mpi_backend():
    MPI.COMM_WORLD.Barrier()


def banner(config, backend, world_size, init_elapsed):
    info = collect_system_info()
    m, inp, ex = config["model"], config["input"], config["execution"]
    lines = [
        "", RULE, f"Experiment: {config['experiment']['name']}", f"Backend: {backend.upper()}", RULE,
        f"World size: {world_size}", f"Initialization time: {init_elapsed:.4f}s", "", "System Info:",
        f"  Platform: {info['platform']}", f"  Python: {info['python_version']}", f"  PyTorch: {info['torch_version']}",
        f"  CPUs (physical): {info['cpu_count']}", f"  CPUs (logical): {info['cpu_count_logical']}",
        f"  Total Memory: {info['total_memory_gb']:.2f} GB",
    ]
    if "gpu_name" in info:
        lines.append(f"  GPUs: {info['gpu_count']} x {info['gpu_name']} ({info['gpu_memory_gb']:.0f} GB, {info['gpu_sm_count']} SMs)")
    lines += ["", f"Model: {m['size']}", f"  Hidden size: {m['hidden_size']}", f"  Layers: {m['num_layers']}",
              f"  Heads: {m['num_heads']}", "", "Input:", f"  Batch size: {inp['batch_size']}",
              f"  Sequence length: {inp['sequence_length']}", "", "Execution:",
              f"  Warmup iterations: {ex['warmup_iterations']}", f"  Benchmark iterations: {ex['benchmark_iterations']}",
              RULE, ""]
    print("\n".join(lines))


def main():
    ap = argparse.ArgumentParser(description="Run the tensor-parallel forward benchmark")
    ap.add_argument("--config", type=str, required=True, help="Path to configuration YAML file")
    ap.add_argument("--backend", type=str, required=True, choices=["openmpi", "intelmpi", "dlbb"],
                    help="backend name (for output file naming)")
    args = ap.parse_args()

    config = load_config(args.config)
    setup_environment(config)

    t0 = time.time()
    rank, world_size, comm = initialize_mpi_backend()
    init_elapsed = time.time() - t0

    expected = config["parallelism"]["world_size"]
    if world_size != expected:
        if rank == 0:
            print(f"ERROR: World size mismatch. Expected {expected}, got {world_size} "
                  f"(set DLBB_WORLD_SIZE={world_size} to override the YAML)")
        sys.exit(1)

    if rank == 0:
        banner(config, args.backend, world_size, init_elapsed)
    comm.Barrier()

    log = (lambda *a: print(f"[Rank {rank}]", *a)) if rank == 0 else (lambda *a: None)
    log("Creating model...")
    model = create_model_from_config(config, rank, world_size, comm)
    log("Model created:")
    log(f"  Total parameters: {model.get_num_parameters() / 1e9:.2f}B")
    log(f"  Memory per rank: {model.get_memory_footprint() / 1e9:.2f} GB")
    comm.Barrier()

    log("Creating dataset...")
    dataset = create_dataset_from_config(config)      # same seed on every rank: replicated input for TP
    batch = dataset.get_batch()
    log("Dataset created:")
    log(f"  Batch shape: {batch.shape}")
    log(f"  Batch dtype: {batch.dtype}")
    log(f"  Batch size: {batch.numel() * batch.element_size() / 1e6:.2f} MB")
    device = model.device
    if device.type == "cuda":
        batch = batch.to(device)                        # resident input: the forward is what is timed
        torch.cuda.synchronize(device)
    comm.Barrier()

    metrics = MetricsCollector(rank, world_size)
    metrics.record_init_time(init_elapsed)

    warmup_iters = config["execution"]["warmup_iterations"]
    log(f"\nStarting warmup: {warmup_iters} iterations")
    for i in range(warmup_iters):
        with Timer() as t:
            model(batch)
            if device.type == "cuda":
                torch.cuda.synchronize(device)
        metrics.record_warmup_time(t.elapsed)
        if (i + 1) % 10 == 0:
            log(f"  Warmup iteration {i + 1}/{warmup_iters}")
    comm.Barrier()

    bench_iters = config["execution"]["benchmark_iterations"]
    log(f"\nStarting benchmark: {bench_iters} iterations")
    for i in range(bench_iters):
        comm.Barrier()
        dt = DeviceTimer(device)
        with Timer() as t:
            with dt:
                model(batch)
            dev_s = dt.sync()                           # forward complete on the device
        comm.Barrier()
        metrics.record_forward_time(t.elapsed, dev_s)
        if (i + 1) % 20 == 0:
            log(f"  Benchmark iteration {i + 1}/{bench_iters}")
    comm.Barrier()

    summary = metrics.get_summary()
    log("\nBenchmark complete. Processing results...")
    per_rank = comm.gather(summary["forward_mean"], root=0)
    per_rank_dev = comm.gather(summary.get("forward_device_mean"), root=0)

    if rank == 0:
        stats = rank_statistics(per_rank)
        if per_rank_dev and per_rank_dev[0] is not None:
            stats["forward_device_mean_per_rank"] = per_rank_dev
            stats["forward_device_max_over_ranks"] = max(per_rank_dev)
        results = {
            "experiment": config["experiment"]["name"],
            "backend": args.backend,
            "config": config,
            "system_info": collect_system_info(),
            "rank_0_summary": summary,
            "rank_statistics": stats,
            "raw_metrics_rank_0": metrics.get_raw_metrics(),
        }
        print_summary(summary, args.backend.upper(), rank)
        print("\nRank Statistics:")
        print(f"  Variance across ranks: {stats['variance_across_ranks']:.6f}")
        print(f"  Coefficient of variation: {stats['coefficient_of_variation']:.4f}")
        print(f"  Forward mean per rank: {stats['forward_mean_per_rank']}")
        out = f"{config['experiment']['output_dir']}/{args.backend}_{config['experiment']['name']}.json"
        save_results(results, out, rank)

    comm.Barrier()
    cleanup_mpi_backend()
    if rank == 0:
        print(f"\n{RULE}\nExperiment completed successfully!\n{RULE}\n")


if __name__ == "__main__":
    main()
