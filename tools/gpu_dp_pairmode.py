#!/usr/bin/env python3
"""DP training step (bench.py's model and trainer) under the three GEMM modes in ONE process (torchrun, N ranks):
0 = one CTA per tile, 1 = CTA pairs, 2 = CTA pairs capped at 168 registers (room for the co-resident bucket kernel)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from distributed_llm_backend_benchmark_b200 import _lib
from distributed_llm_backend_benchmark_b200.utils.config import load_config
from distributed_llm_backend_benchmark_b200.train import Trainer

world, rank, lr = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
if world > 1: dist.init_process_group("nccl", device_id=dev)
cfg = load_config(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "config", "baseline_config.yaml"))
trainer = Trainer(cfg, device=dev, rank=rank, world_size=world, backend="native")
x, t = trainer.resident_batch()
lib = _lib.load()


def barrier():
    if world > 1: dist.barrier()
    torch.cuda.synchronize()


rows = []
steps = int(os.environ.get("STEPS", "6"))
for rnd in range(2):
    for mode in (1, 0, 2):
        lib.dlbb_gemm_set_pair_mode(mode)
        for _ in range(2): trainer.step(x, t)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps): loss = trainer.step(x, t)
        e1.record(); barrier()
        ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
        if world > 1: dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        row = {"world": world, "round": rnd, "pair_mode": mode, "ms_per_step": round(ms.item(), 2), "loss": float(loss.item())}
        rows.append(row)
        if rank == 0: print("PAIRMODE", json.dumps(row), flush=True)
lib.dlbb_gemm_set_pair_mode(-1)
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True); json.dump(rows, open(f"gpurun_out/dp_pairmode_n{world}.json", "w"), indent=1)
if world > 1: dist.barrier(); dist.destroy_process_group()
